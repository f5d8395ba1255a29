// oracle/powf_check.cpp — TEST INFRASTRUCTURE.  Compares prt_b200/csrc/prt_powf.h (the restatement the device runs)
// with the host libm's powf, bit for bit: every fp32 in [0, 1] (and a band above 1) for y = 2.2f and y = 1/2.2f — the only
// exponents the reference uses (material.cpp:24-28, image.cpp:16-19) — plus random (x, y) pairs.  Prints mismatch counts.
//   g++ -O2 -ffp-contract=off powf_check.cpp -o powf_check && ./powf_check
#include <cstdio>
#include <cstdint>
#include <cmath>
#include <thread>
#include <vector>
#include <atomic>
#include "../prt_b200/csrc/prt_powf.h"

int main() {
    const float ys[2] = {2.2f, 1 / 2.2f};
    std::atomic<uint64_t> bad{0}, total{0};
    const uint32_t hi = prt_powf::asU32(1.5f);
    const int T = (int)std::thread::hardware_concurrency() ? (int)std::thread::hardware_concurrency() : 4;
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++) th.emplace_back([&, t]() {
        uint64_t b = 0, n = 0;
        for (uint32_t u = (uint32_t)t; u <= hi; u += (uint32_t)T)
            for (float y : ys) {
                float x = prt_powf::asF32(u);
                volatile float vx = x, vy = y;
                float a = powf(vx, vy), r = prt_powf::powfRef(x, y);
                n++;
                if (prt_powf::asU32(a) != prt_powf::asU32(r)) { if (b < 3) std::printf("x=%a y=%a libm=%a ours=%a\n", x, y, a, r); b++; }
            }
        uint32_t s = 12345u + 977u * (uint32_t)t;
        auto rnd = [&]() { s ^= s << 13; s ^= s >> 17; s ^= s << 5; return s; };
        for (int i = 0; i < 2000000; i++) {
            float x = prt_powf::asF32(rnd() & 0x7fffffffu), y = ((int)(rnd() % 20001) - 10000) / 997.0f + 0.0001234f;
            if (!(x == x) || y == std::floor(y)) continue;
            volatile float vx = x, vy = y;
            float a = powf(vx, vy), r = prt_powf::powfRef(x, y);
            n++;
            if (prt_powf::asU32(a) != prt_powf::asU32(r) && !(a != a && r != r)) { if (b < 3) std::printf("x=%a y=%a libm=%a ours=%a\n", x, y, a, r); b++; }
        }
        bad += b; total += n;
    });
    for (auto& x : th) x.join();
    std::printf("compared %llu arguments: %llu mismatches\n", (unsigned long long)total.load(), (unsigned long long)bad.load());
    return bad.load() ? 1 : 0;
}
