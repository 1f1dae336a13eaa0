// temporary: replaced by rnnoise.cu
#include "rnnoise.cuh"
namespace fvad {
int rnnoise_batch_create(size_t, size_t, int, RnnoiseBatch **out) { *out = nullptr; return set_error(FVAD_EUNSUPPORTED, "denoiser not built yet"); }
void rnnoise_batch_destroy(RnnoiseBatch *) {}
int rnnoise_batch_reset(RnnoiseBatch *, cudaStream_t) { return FVAD_OK; }
int rnnoise_batch_process(RnnoiseBatch *, const float *, size_t, int, size_t, float *, size_t, size_t, float *, float *, size_t, size_t, cudaStream_t) { return FVAD_EUNSUPPORTED; }
}
struct fvad_denoiser; struct fvad_denoiser_batch;
extern "C" {
int fvad_denoiser_create(fvad_denoiser **) { return fvad::set_error(FVAD_EUNSUPPORTED, "denoiser not built yet"); }
void fvad_denoiser_destroy(fvad_denoiser *) {}
size_t fvad_denoiser_frame_size(void) { return 480; }
int fvad_denoiser_denoise(fvad_denoiser *, const float *, size_t, const float *, size_t, float *, size_t, float *) { return FVAD_EUNSUPPORTED; }
int fvad_denoiser_batch_create(size_t, int, fvad_denoiser_batch **) { return FVAD_EUNSUPPORTED; }
void fvad_denoiser_batch_destroy(fvad_denoiser_batch *) {}
int fvad_denoiser_batch_run(fvad_denoiser_batch *, const float *const *, size_t, float *const *, float *const *) { return FVAD_EUNSUPPORTED; }
}
