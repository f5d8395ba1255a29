// oracle/ref_p1_shim.h — force-included (-include) into every reference translation unit.  TEST INFRASTRUCTURE.
//
// SURVEY §8c patch P1, applied WITHOUT touching the reference sources: clang accepts the implicit
// __m256i -> __m256 conversions at vecmath.h:332 (SoaMask(int32_t)) and vecmath.h:384 (SoaMask::computeNot);
// g++ does not.  Those are the only two uses of these intrinsics in the reference, so wrapping the
// intrinsics themselves in a bit-cast is exactly the "two bit-neutral casts" of the survey.
#pragma once
#include <immintrin.h>
#define _mm256_set_epi32(...) _mm256_castsi256_ps((_mm256_set_epi32)(__VA_ARGS__))
#define _mm256_cmpeq_epi32(a, b) _mm256_castsi256_ps((_mm256_cmpeq_epi32)((a), (b)))
