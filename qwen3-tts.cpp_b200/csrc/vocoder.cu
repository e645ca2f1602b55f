// placeholder until the vocoder kernels land (next commit)
#include "common.cuh"
#include "engine.h"
namespace q3t {
int  vocoder_finalize(q3t_ctx * ctx) { return set_err(ctx, Q3T_ERR_STATE, "vocoder not built yet"); }
void vocoder_free(q3t_ctx *) {}
int64_t vocoder_run(q3t_ctx * ctx, const int32_t *, int, float **) { return set_err(ctx, Q3T_ERR_STATE, "vocoder not built yet"); }
}
extern "C" int64_t q3t_vocoder_n_samples(int32_t n_frames) { return n_frames > 0 ? 1920LL * n_frames - 555 : 0; }
extern "C" int64_t q3t_vocoder_decode(q3t_ctx * ctx, const int32_t *, int32_t, float *, int64_t) { return q3t::set_err(ctx, Q3T_ERR_STATE, "vocoder not built yet"); }
extern "C" int64_t q3t_fetch_samples(q3t_ctx * ctx, float *, int64_t) { return q3t::set_err(ctx, Q3T_ERR_STATE, "vocoder not built yet"); }
extern "C" int q3t_codepred_parallel(q3t_ctx * ctx, const float *, const int32_t *, float *) { return q3t::set_err(ctx, Q3T_ERR_STATE, "not built yet"); }
