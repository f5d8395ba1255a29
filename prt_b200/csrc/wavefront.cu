// placeholder until the wavefront path tracer lands (next commit)
#include "prt_context.h"
using namespace prt_b200;
namespace prt_b200 { void destroyWavefront(prt_ctx*) {} }
extern "C" {
int prt_cuda_render_dev(prt_ctx* ctx, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, float, float*) { return setError(ctx, PRT_ERR_STATE, "render: not built yet"); }
int prt_cuda_render(prt_ctx* ctx, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, float, float*) { return setError(ctx, PRT_ERR_STATE, "render: not built yet"); }
}
