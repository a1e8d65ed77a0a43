// Tracking multicorrelator engine: C ABI (include/gnss_b200.h) over trk_kernel.cuh.
#include "trk_kernel2.cuh"
#include "trk_kernel3.cuh"
#include "trk_loop.cuh"
#include "ingest_kernel.cuh"
#include "host_pool.h"
#include <chrono>
#include <mutex>
#include <algorithm>
#include <cmath>
#include <complex>
#include <cstring>
#include <new>
#include <vector>

using namespace gb200;

struct gb200_trk
{
    gb200_trk_config cfg{};
    cudaStream_t stream{};
    cudaEvent_t ev0{}, ev1{};
    int elem_bytes{8};
    // IQ ring: capacity + max_block samples; the first max_block samples are mirrored behind the end so that every
    // window of up to max_block samples is contiguous
    char* d_ring{nullptr};
    uint64_t head{0};  // absolute index one past the newest sample
    // code tables, one slot of max_code_len float2 per channel (real tables use the first half)
    float* d_codes{nullptr};
    std::vector<int> code_len;
    std::vector<char> code_cplx;
    // command staging
    TrkDevCmd* h_cmds{nullptr};  // pinned
    TrkDevCmd* d_cmds{nullptr};
    float* d_out{nullptr};
    float* h_out{nullptr};  // pinned
    float* d_partials{nullptr};
    unsigned* d_counters{nullptr};
    size_t partial_ctas{0};
    bool scan_ok{false};                            // all commands of the call pass run_applicable (set with the preprocessing)
    float scan_min_step{0}, scan_max_step{0};
    int last_path{0}, last_R{18};  // path 0: chunked kernel, 1: run-based kernel, 2: high-dynamics kernel
    // last enqueued launch (for collect / replay)
    int last_n{0}, last_T{0}, last_chunks{0}, last_cplx{0};
    size_t last_smem{0};
    int sm_count{148};
    int max_smem{0};
    unsigned char* d_raw{nullptr};  // staging of raw ingest bytes (gb200_trk_push_raw)
    size_t raw_bytes{0};
};

namespace
{
constexpr double TWO_PI = 6.283185307179586476925286766559;
constexpr int RING_SLACK = 4096;

// samples per run of the run-based kernel for a sample type (long / short variant)
constexpr int trk_run_len(int iq, bool long_run)
{
    return iq == GB200_IQ_CF32 ? (long_run ? 18 : 10) : iq == GB200_IQ_CS16 ? (long_run ? 20 : 12) : (long_run ? 24 : 8);
}
// extra-long variant (cf32 only: 22 samples = 11 chunks; the next odd chunk count of the integer types is already past one table entry
// per run at every rate where their long variant applies).  Measured at 10 240 channels x 25 000 samples x 11 taps: 0.295 against
// 0.306 ms per step (fewer runs per sample: the per-run boundary search and the tap phase are amortised over 22 samples)
constexpr int TRK_RUN_XLONG = 22;
// one table boundary per run at most: the law's argument grows by (R - 1) * step over a run, kept below 1 with room for its fp32 rounding
inline bool trk_run_xlong_fits(int iq, double max_step) { return iq == GB200_IQ_CF32 && max_step * (TRK_RUN_XLONG - 1) <= 0.92; }

// fraction of a turn in [0,1) -> 64-bit fixed point
inline unsigned long long turns_to_fix(double turns)
{
    double f = turns - std::floor(turns);
    if (f >= 1.0) f = 0.0;
    return (unsigned long long)std::ldexp(f, 64 - 11) << 11;  // 53 significant bits, shifted to the top
}

template <int T, int IQ, bool CPLX>
cudaError_t launch_one(gb200_trk* h, int n_cmds, int n_chunks, size_t smem)
{
    if (smem > 48 * 1024)
        {
            cudaError_t e = cudaFuncSetAttribute(trk_correlate_kernel<T, IQ, CPLX>, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem);
            if (e != cudaSuccess) return e;
        }
    dim3 grid(n_chunks, n_cmds);
    trk_correlate_kernel<T, IQ, CPLX><<<grid, TRK_THREADS, smem, h->stream>>>(h->d_cmds, n_chunks, h->d_partials, h->d_counters, h->d_out);
    count_launch();
    return cudaGetLastError();
}

template <int IQ, bool CPLX>
cudaError_t launch_T(gb200_trk* h, int T, int n_cmds, int n_chunks, size_t smem)
{
    switch (T)
        {
        case 1: return launch_one<1, IQ, CPLX>(h, n_cmds, n_chunks, smem);
        case 2: return launch_one<2, IQ, CPLX>(h, n_cmds, n_chunks, smem);
        case 3: return launch_one<3, IQ, CPLX>(h, n_cmds, n_chunks, smem);
        case 5: return launch_one<5, IQ, CPLX>(h, n_cmds, n_chunks, smem);
        case 8: return launch_one<8, IQ, CPLX>(h, n_cmds, n_chunks, smem);
        case 11: return launch_one<11, IQ, CPLX>(h, n_cmds, n_chunks, smem);
        default: return launch_one<16, IQ, CPLX>(h, n_cmds, n_chunks, smem);
        }
}

int padded_taps(int t)
{
    static const int sizes[] = {1, 2, 3, 5, 8, 11, 16};
    for (int s : sizes)
        if (t <= s) return s;
    return 16;
}

template <int T, int IQ, int CH>
cudaError_t launch_run_one(gb200_trk* h)
{
    const size_t smem = trk2_smem_bytes<IQ, CH>();
    cudaError_t e = cudaFuncSetAttribute(trk_run_kernel<T, IQ, CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    trk_run_kernel<T, IQ, CH><<<h->last_n, TRK2_THREADS, smem, h->stream>>>(h->d_cmds, reinterpret_cast<const unsigned char*>(h->d_ring), h->d_out);
    count_launch();
    return cudaGetLastError();
}

// run lengths: an odd number of 16-byte chunks (see trk_kernel2.cuh); long and short variant per sample type
template <int T, int IQ>
cudaError_t launch_run_RG(gb200_trk* h)
{
    constexpr int CH_LONG = IQ == GB200_IQ_CF32 ? 9 : IQ == GB200_IQ_CS16 ? 5 : 3;
    constexpr int CH_SHORT = IQ == GB200_IQ_CF32 ? 5 : IQ == GB200_IQ_CS16 ? 3 : 1;
    if (IQ == GB200_IQ_CF32 && h->last_R == TRK_RUN_XLONG) return launch_run_one<T, GB200_IQ_CF32, 11>(h);
    return h->last_R == trk_run_len(IQ, true) ? launch_run_one<T, IQ, CH_LONG>(h) : launch_run_one<T, IQ, CH_SHORT>(h);
}

template <int IQ>
cudaError_t launch_run_T(gb200_trk* h)
{
    switch (h->last_T)
        {
        case 3: return launch_run_RG<3, IQ>(h);
        case 5: return launch_run_RG<5, IQ>(h);
        default: return launch_run_RG<11, IQ>(h);
        }
}

cudaError_t launch_run(gb200_trk* h)
{
    switch (h->cfg.iq_dtype)
        {
        case GB200_IQ_CF32: return launch_run_T<GB200_IQ_CF32>(h);
        case GB200_IQ_CS16: return launch_run_T<GB200_IQ_CS16>(h);
        default: return launch_run_T<GB200_IQ_CS8>(h);
        }
}

// Where the carrier angles of a command are computed: on the device by trk_phase_kernel (default) or with the host's float64
// atan2 (GB200_HOST_PHASES=1; 50 ns more host time per command).  The two differ by rounding in the last bits of a 64-bit angle.
constexpr bool kDevicePhasesDefault = true;
inline bool device_phases()
{
    static const bool on = [] {
        const char* e = getenv("GB200_HOST_PHASES");
        return e ? std::atoi(e) == 0 : kDevicePhasesDefault;
    }();
    return on;
}

// phasors -> fixed-point angles, in place in d_cmds (trk_phase_kernel); queued behind the command upload on the engine's stream
cudaError_t upload_phases(gb200_trk* h, int n_cmds)
{
    if (!device_phases()) return cudaSuccess;
    trk_phase_kernel<<<(n_cmds + 255) / 256, 256, 0, h->stream>>>(h->d_cmds, n_cmds);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch(gb200_trk* h)
{
    if (h->last_path == 1) return launch_run(h);
    if (h->last_path == 2)
        {
            switch (h->cfg.iq_dtype)
                {
                case GB200_IQ_CF32: trk_high_dyn_kernel<GB200_IQ_CF32><<<h->last_n, TRK3_THREADS, 0, h->stream>>>(h->d_cmds, h->d_out); break;
                case GB200_IQ_CS16: trk_high_dyn_kernel<GB200_IQ_CS16><<<h->last_n, TRK3_THREADS, 0, h->stream>>>(h->d_cmds, h->d_out); break;
                default: trk_high_dyn_kernel<GB200_IQ_CS8><<<h->last_n, TRK3_THREADS, 0, h->stream>>>(h->d_cmds, h->d_out); break;
                }
            count_launch();
            return cudaGetLastError();
        }
    const int T = h->last_T, n = h->last_n, ch = h->last_chunks;
    const size_t sm = h->last_smem;
    if (h->last_cplx)
        {
            switch (h->cfg.iq_dtype)
                {
                case GB200_IQ_CF32: return launch_T<GB200_IQ_CF32, true>(h, T, n, ch, sm);
                case GB200_IQ_CS16: return launch_T<GB200_IQ_CS16, true>(h, T, n, ch, sm);
                default: return launch_T<GB200_IQ_CS8, true>(h, T, n, ch, sm);
                }
        }
    switch (h->cfg.iq_dtype)
        {
        case GB200_IQ_CF32: return launch_T<GB200_IQ_CF32, false>(h, T, n, ch, sm);
        case GB200_IQ_CS16: return launch_T<GB200_IQ_CS16, false>(h, T, n, ch, sm);
        default: return launch_T<GB200_IQ_CS8, false>(h, T, n, ch, sm);
        }
}
}  // namespace

extern "C" {

int gb200_trk_create(const gb200_trk_config* cfg, gb200_trk** out)
{
    if (!cfg || !out) return set_error(GB200_EINVAL, "null argument");
    *out = nullptr;
    if (cfg->max_channels <= 0 || cfg->max_code_len <= 0 || cfg->max_block <= 0 || cfg->max_cmds <= 0)
        return set_error(GB200_EINVAL, "max_channels, max_code_len, max_block and max_cmds must be positive");
    if (cfg->iq_dtype < GB200_IQ_CF32 || cfg->iq_dtype > GB200_IQ_CS8) return set_error(GB200_EINVAL, "unknown iq_dtype %d", cfg->iq_dtype);
    if (cfg->ring_samples < 2LL * cfg->max_block) return set_error(GB200_EINVAL, "ring_samples must be >= 2*max_block");
    if (cfg->max_block > (1 << 22)) return set_error(GB200_EINVAL, "max_block above 2^22 samples is not supported");
    if (cfg->max_cmds > 65535) return set_error(GB200_EINVAL, "max_cmds %d above 65535 (one launch indexes its commands with gridDim.y)", cfg->max_cmds);
    int rc = select_device(cfg->device);
    if (rc) return rc;
    gb200_trk* h = new (std::nothrow) gb200_trk;
    if (!h) return set_error(GB200_ENOMEM, "out of host memory");
    // every allocation below goes through one failure path: a handle that could not be completed is torn down, never leaked
    rc = [&]() -> int {
    h->cfg = *cfg;
    h->elem_bytes = cfg->iq_dtype == GB200_IQ_CF32 ? 8 : cfg->iq_dtype == GB200_IQ_CS16 ? 4 : 2;
    cudaDeviceProp p;
    GB_CUDA(cudaGetDeviceProperties(&p, cfg->device));
    h->sm_count = p.multiProcessorCount;
    h->max_smem = (int)p.sharedMemPerBlockOptin - 2048;
    GB_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
    GB_CUDA(cudaEventCreate(&h->ev0));
    GB_CUDA(cudaEventCreate(&h->ev1));
    // + RING_SLACK: the run-based kernel stages whole tiles, which may extend past the last window (masked, but must be mapped)
    GB_CUDA(cudaMalloc(&h->d_ring, (size_t)(cfg->ring_samples + cfg->max_block + RING_SLACK) * h->elem_bytes));
    GB_CUDA(cudaMemsetAsync(h->d_ring, 0, (size_t)(cfg->ring_samples + cfg->max_block + RING_SLACK) * h->elem_bytes, h->stream));
    GB_CUDA(cudaMalloc(&h->d_codes, (size_t)cfg->max_channels * cfg->max_code_len * sizeof(float2)));
    h->code_len.assign(cfg->max_channels, 0);
    h->code_cplx.assign(cfg->max_channels, 0);
    GB_CUDA(cudaMallocHost(&h->h_cmds, (size_t)cfg->max_cmds * sizeof(TrkDevCmd)));
    GB_CUDA(cudaMalloc(&h->d_cmds, (size_t)cfg->max_cmds * sizeof(TrkDevCmd)));
    GB_CUDA(cudaMalloc(&h->d_out, (size_t)cfg->max_cmds * 2 * TRK_OUT_STRIDE * sizeof(float)));
    GB_CUDA(cudaMallocHost(&h->h_out, (size_t)cfg->max_cmds * 2 * TRK_OUT_STRIDE * sizeof(float)));
    const size_t min_chunks = (size_t)(cfg->max_block + 16383) / 16384 + 1;
    h->partial_ctas = std::max<size_t>(16 * (size_t)h->sm_count, (size_t)cfg->max_cmds * min_chunks);
    GB_CUDA(cudaMalloc(&h->d_partials, h->partial_ctas * 2 * TRK_OUT_STRIDE * sizeof(float)));
    GB_CUDA(cudaMalloc(&h->d_counters, (size_t)cfg->max_cmds * sizeof(unsigned)));
    GB_CUDA(cudaMemsetAsync(h->d_counters, 0, (size_t)cfg->max_cmds * sizeof(unsigned), h->stream));
    GB_CUDA(cudaStreamSynchronize(h->stream));
    return GB200_OK;
    }();
    if (rc)
        {
            gb200_trk_destroy(h);
            return rc;
        }
    *out = h;
    return GB200_OK;
}

int gb200_trk_destroy(gb200_trk* h)
{
    if (!h) return GB200_OK;
    cudaSetDevice(h->cfg.device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    cudaFree(h->d_ring);
    cudaFree(h->d_raw);
    cudaFree(h->d_codes);
    cudaFreeHost(h->h_cmds);
    cudaFree(h->d_cmds);
    cudaFree(h->d_out);
    cudaFreeHost(h->h_out);
    cudaFree(h->d_partials);
    cudaFree(h->d_counters);
    if (h->ev0) cudaEventDestroy(h->ev0);
    if (h->ev1) cudaEventDestroy(h->ev1);
    if (h->stream) cudaStreamDestroy(h->stream);
    (void)cudaGetLastError();
    delete h;
    return GB200_OK;
}

static int set_code_impl(gb200_trk* h, int ch, const void* code, int code_len, bool cplx)
{
    if (!h || !code) return set_error(GB200_EINVAL, "null argument");
    if (ch < 0 || ch >= h->cfg.max_channels) return set_error(GB200_EINVAL, "channel %d out of range", ch);
    if (code_len <= 0 || code_len > h->cfg.max_code_len) return set_error(GB200_EINVAL, "code_len %d outside 1..%d", code_len, h->cfg.max_code_len);
    GB_CUDA(cudaSetDevice(h->cfg.device));
    float* dst = h->d_codes + (size_t)ch * h->cfg.max_code_len * 2;
    GB_CUDA(cudaMemcpyAsync(dst, code, (size_t)code_len * (cplx ? sizeof(float2) : sizeof(float)), cudaMemcpyHostToDevice, h->stream));
    GB_CUDA(cudaStreamSynchronize(h->stream));  // the caller's buffer may be reused on return
    h->code_len[ch] = code_len;
    h->code_cplx[ch] = cplx;
    return GB200_OK;
}

int gb200_trk_set_code(gb200_trk* h, int ch, const float* code, int code_len) { return set_code_impl(h, ch, code, code_len, false); }
int gb200_trk_set_code_complex(gb200_trk* h, int ch, const gb200_cf32* code, int code_len) { return set_code_impl(h, ch, code, code_len, true); }

// copy n samples to the ring positions of absolute indices [abs_start, abs_start + n), with the mirror of [0, max_block)
static int write_at(gb200_trk* h, uint64_t abs_start, const void* src, int64_t n, cudaMemcpyKind kind, cudaStream_t stream = nullptr)
{
    if (!stream) stream = h->stream;
    const int64_t cap = h->cfg.ring_samples, mb = h->cfg.max_block;
    const int eb = h->elem_bytes;
    int64_t done = 0;
    while (done < n)
        {
            const int64_t pos = (int64_t)((abs_start + (uint64_t)done) % (uint64_t)cap);
            const int64_t run = std::min<int64_t>(n - done, cap - pos);
            const char* s = static_cast<const char*>(src) + done * eb;
            GB_CUDA(cudaMemcpyAsync(h->d_ring + pos * eb, s, (size_t)run * eb, kind, stream));
            if (pos < mb)  // mirror the part that falls in [0, max_block) behind the end of the ring
                {
                    const int64_t m = std::min<int64_t>(run, mb - pos);
                    GB_CUDA(cudaMemcpyAsync(h->d_ring + (cap + pos) * eb, s, (size_t)m * eb, kind, stream));
                }
            done += run;
        }
    return GB200_OK;
}

static int push_impl(gb200_trk* h, const void* src, int64_t n, uint64_t* first_index, cudaMemcpyKind kind, cudaStream_t stream = nullptr)
{
    if (!h || (!src && n > 0)) return set_error(GB200_EINVAL, "null argument");
    if (n < 0 || n > h->cfg.ring_samples) return set_error(GB200_EINVAL, "push of %lld samples exceeds the ring (%lld)", (long long)n, (long long)h->cfg.ring_samples);
    GB_CUDA(cudaSetDevice(h->cfg.device));
    if (first_index) *first_index = h->head;
    const int rc = write_at(h, h->head, src, n, kind, stream);
    if (rc) return rc;
    h->head += (uint64_t)n;
    return GB200_OK;
}

int gb200_trk_push_iq(gb200_trk* h, const void* iq_host, int64_t n_samples, uint64_t* first_index)
{
    int rc = push_impl(h, iq_host, n_samples, first_index, cudaMemcpyHostToDevice);
    if (rc) return rc;
    GB_CUDA(cudaStreamSynchronize(h->stream));  // pageable or pinned: the host buffer is free on return
    return GB200_OK;
}

int gb200_trk_push_iq_device(gb200_trk* h, const void* iq_dev, int64_t n_samples, uint64_t* first_index)
{
    return push_impl(h, iq_dev, n_samples, first_index, cudaMemcpyDeviceToDevice);
}

int gb200_trk_skip(gb200_trk* h, int64_t n_samples)
{
    if (!h || n_samples < 0) return set_error(GB200_EINVAL, "bad argument");
    h->head += (uint64_t)n_samples;
    return GB200_OK;
}

int gb200_trk_write_iq(gb200_trk* h, uint64_t abs_index, const void* iq_host, int64_t n_samples)
{
    if (!h || (!iq_host && n_samples > 0) || n_samples < 0) return set_error(GB200_EINVAL, "bad argument");
    const uint64_t cap = (uint64_t)h->cfg.ring_samples;
    if (abs_index + (uint64_t)n_samples > h->head) return set_error(GB200_ERANGE, "write [%llu,+%lld) beyond the newest sample %llu", (unsigned long long)abs_index, (long long)n_samples, (unsigned long long)h->head);
    if (h->head > cap && abs_index < h->head - cap) return set_error(GB200_ERANGE, "write start %llu already overwritten (ring holds the last %llu samples)", (unsigned long long)abs_index, (unsigned long long)cap);
    GB_CUDA(cudaSetDevice(h->cfg.device));
    const int rc = write_at(h, abs_index, iq_host, n_samples, cudaMemcpyHostToDevice);
    if (rc) return rc;
    GB_CUDA(cudaStreamSynchronize(h->stream));
    return GB200_OK;
}

uint64_t gb200_trk_head(const gb200_trk* h) { return h ? h->head : 0; }


}  // extern "C"

#include "ingest.inc"
#include "trk_submit.inc"
#include "trk_loop.inc"
#include "trk_noalloc.inc"
