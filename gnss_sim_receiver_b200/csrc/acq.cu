// PCPS acquisition engine — placeholder entry points until the FFT kernels land (next commit).
#include "common.cuh"
using namespace gb200;
#define NOTYET return set_error(GB200_ESTATE, "%s: acquisition engine not built yet", __func__)
extern "C" {
int gb200_acq_create(const gb200_acq_config*, gb200_acq**) { NOTYET; }
int gb200_acq_destroy(gb200_acq*) { return GB200_OK; }
int gb200_acq_set_local_code(gb200_acq*, int, const gb200_cf32*, int) { NOTYET; }
int gb200_acq_set_doppler_grid(gb200_acq*, int, int, int) { NOTYET; }
int gb200_acq_num_doppler_bins(const gb200_acq*) { return 0; }
int gb200_acq_neff(const gb200_acq*) { return 0; }
int gb200_acq_run(gb200_acq*, const void*, int, const int32_t*, int, gb200_acq_result*) { NOTYET; }
int gb200_acq_run_device(gb200_acq*, const void*, int, const int32_t*, int, gb200_acq_result*) { NOTYET; }
int gb200_acq_replay(gb200_acq*, int, float*) { NOTYET; }
float gb200_acq_threshold(const gb200_acq*, float) { return 0.0f; }
int gb200_acq_keep_grid(gb200_acq*, int) { NOTYET; }
int gb200_acq_read_grid(gb200_acq*, int, float*) { NOTYET; }
int gb200_acq_fft(gb200_acq*, const gb200_cf32*, gb200_cf32*, int) { NOTYET; }
}
