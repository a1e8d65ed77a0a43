// PCPS acquisition engine: C ABI (include/gnss_b200.h) over acq_kernel.cuh.
#include "acq_kernel.cuh"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <new>
#include <vector>

using namespace gb200;

struct gb200_acq
{
    gb200_acq_config cfg{};
    cudaStream_t stream{};
    cudaEvent_t ev0{}, ev1{};
    FftPlan plan{};
    int elem_bytes{8};
    int threads{1024};
    size_t smem{0};
    int n_bins{0}, neff{0}, out_offset{0}, n_parts{1};
    // transform length: cfg.fft_size, or — for sizes the mixed-radix plan cannot factor (a prime factor above 61) — a smooth length
    // >= 2 * fft_size - 1 on which the same circular correlation is computed exactly (zero-padded input, wrapped replica, search over
    // the first fft_size lags only): FFTW takes any size, this is how the engine does
    int tlen{0};
    bool padded{false};
    bool cluster_inv{false};  // split transforms: the blocks of a cell run as a thread-block cluster (acq_inv_kernel<3>) instead of staging in global memory
    int doppler_max{0}, doppler_step{0}, doppler_center{0};
    int keep_grid{0}, user_keep{0};
    std::vector<int> pos_of_k;  // natural index -> digit-reversed position
    float2 *d_twA{nullptr}, *d_twB{nullptr};  // two-level W_N table for the global stages
    float2* d_twT{nullptr};                   // per-stage tables of the shared-memory stages (FftPlan::tw_off)
    std::vector<float2> tw_stage;
    // forward transforms of the wiped input may run as (global first factor) + (several small shared-memory blocks per bin):
    // same digit order as `plan`, but 4-5 times as many CTAs, so all SMs work on the n_bins forward transforms
    FftPlan plan_f{};
    bool split_fwd{false};
    int threads_f{0};
    size_t smem_f{0};
    float2* d_twT_f{nullptr};
    std::vector<float2> tw_stage_f;
    void* d_iq{nullptr};
    float2* d_X{nullptr};        // [n_bins][N] forward spectra of the wiped input
    float2* d_C{nullptr};        // [max_prns][N] conjugated code spectra
    float2* d_stage{nullptr};    // [max(n_bins,1)][N] global-stage staging for forward transforms
    float2* d_scratch{nullptr};  // [cells][N] inverse global-stage staging
    size_t scratch_cells{0};
    float* d_grid{nullptr};      // [max_prns][n_bins][neff]
    AcqCell* d_cells{nullptr};
    int* d_slots{nullptr};
    unsigned long long* d_dphi{nullptr};
    float* d_phase{nullptr};     // [n_bins][consumed] fp32 accumulated wipe-off phases (reference law); unused in exact mode
    float* d_inc{nullptr};
    int exact_wipeoff{0};
    gb200_acq_result* d_results{nullptr};
    gb200_acq_result* h_results{nullptr};
    std::vector<char> code_set;
    std::vector<int> dwells_done;  // per slot: dwells accumulated in the grid
    // folded input (QuickSync, pcps_quicksync_acquisition_cc.cc:296-318): a block of fold_len samples is wiped off over its whole length and
    // its first fold_chunks * fft_size samples are added up chunk by chunk into ONE fft_size-point transform input per Doppler bin
    int fold_chunks{0}, fold_len{0}, fold_bins{0};
    float2* d_fold_iq{nullptr};   // [fold_len] the block as uploaded
    float* d_fold_phase{nullptr};  // [n_bins][fold_len] fp32 accumulated wipe-off phases over the whole block
    float2* d_fold{nullptr};       // [n_bins][N] folded, wiped input of every bin
    int fold_alloc_bins{0};        // bins the two tables above are allocated for (kept across Doppler-grid re-installs of the same size)
    // scratch of the strategy helpers (spectrum peak, grid pair maximum / probe, folded candidates): grown on demand, never shrunk
    char* d_aux{nullptr};
    size_t aux_bytes{0};
    // last run (replay)
    std::vector<int> last_slots;
    std::vector<int> dev_slots;  // what d_slots holds
    int last_dwell{0};
    bool have_run{false};
};

namespace
{
constexpr double TWO_PI = 6.283185307179586476925286766559;

// Stage radices in DIF order, even factors first (the late, short-stride stages are the odd ones).  Pairs of equal small
// factors are fused into one register-resident pass (16 = 4x4, 8 = 4x2, 9 = 3x3, 25 = 5x5; acq_fft.cuh: fft_smem_stage2):
// 25000 = 8 * 25 * 25 * 5 takes 4 passes over shared memory instead of 7.  GB200_ACQ_PLAIN_RADIX=1 keeps the plain stages.
bool is_small_radix(int r) { return r == 2 || r == 3 || r == 4 || r == 5 || r == 7 || r == 8 || r == 9 || r == 10 || r == 16 || r == 20 || r == 25 || r == 32; }

bool factor_235_impl(int n, std::vector<int>& radix, bool fuse, bool allow32)
{
    radix.clear();
    const int n_in = n;
    int c2 = 0, c3 = 0, c5 = 0, c7 = 0;
    while (n % 2 == 0) { c2++; n /= 2; }
    while (n % 3 == 0) { c3++; n /= 3; }
    while (n % 5 == 0) { c5++; n /= 5; }
    while (n % 7 == 0) { c7++; n /= 7; }
    // what is left must split into primes 11 ... 61 (generic O(P) per point stage, acq_fft.cuh: fft_smem_stage_prime)
    std::vector<int> primes;
    for (int p = 11; p <= 61 && n > 1; p += 2)
        while (n % p == 0) { primes.push_back(p); n /= p; }
    if (n != 1) return false;
    if (fuse)
        {
            // 2^5 as ONE radix-32 stage (4 x 8 in registers) when it saves a pass: five or seven twos would otherwise cost 16 * 2 or 16 * 8
            // (64 data registers per thread: no spills in the inverse kernels at the 128-register cap; GB200_ACQ_NO_RADIX32 switches it off)
            if ((c2 == 5 || c2 == 7) && allow32)
                {
                    radix.push_back(32);
                    c2 -= 5;
                }
            while (c2 >= 4) { radix.push_back(16); c2 -= 4; }
            if (c2 == 3) { radix.push_back(8); c2 = 0; }
        }
    // a stray 4 or 2 next to an ODD number of factors 5: fused with one of them (radix 20 / 10) the pair costs one pass instead of two and the
    // remaining fives pair up into radix-25 stages (4000 = 16 * 10 * 25, 8000 = 16 * 20 * 25: three passes instead of four)
    if (fuse && (c5 & 1) && getenv("GB200_ACQ_NO_RADIX10") == nullptr)
        {
            if (c2 == 2) { radix.push_back(20); c2 = 0; c5--; }
            else if (c2 == 1) { radix.push_back(10); c2 = 0; c5--; }
        }
    while (c2 >= 2) { radix.push_back(4); c2 -= 2; }
    if (c2) radix.push_back(2);
    if (fuse)
        while (c3 >= 2) { radix.push_back(9); c3 -= 2; }
    while (c3--) radix.push_back(3);
    if (fuse)
        while (c5 >= 2) { radix.push_back(25); c5 -= 2; }
    while (c5--) radix.push_back(5);
    while (c7--) radix.push_back(7);
    if (!primes.empty())
        {
            // the first and the last stage carry the hooks of the inverse kernel: the prime stages go in between
            if (radix.size() < 2) return fuse ? factor_235_impl(n_in, radix, false, false) : false;  // 16 -> 4 * 4 makes room
            radix.insert(radix.begin() + 1, primes.begin(), primes.end());
        }
    return (int)radix.size() <= FFT_MAX_STAGES;
}

// the factorisation with the fewest passes; the radix-32 stage is taken only when it saves one (at equal count 16 * 10 * 25 beats 32 * 25 * 5:
// C1 sweep 0.096 against 0.102 ms)
bool factor_235(int n, std::vector<int>& radix, bool fuse)
{
    if (!factor_235_impl(n, radix, fuse, false)) return false;
    if (fuse && getenv("GB200_ACQ_NO_RADIX32") == nullptr)
        {
            std::vector<int> with32;
            if (factor_235_impl(n, with32, fuse, true) && with32.size() < radix.size()) radix = with32;
        }
    return true;
}

// the plain factors of a (possibly fused) stage radix, in the order the stage applies them
void stage_factors(int r, int* f1, int* f2)
{
    switch (r)
        {
        case 16: *f1 = 4; *f2 = 4; break;
        case 8: *f1 = 4; *f2 = 2; break;
        case 9: *f1 = 3; *f2 = 3; break;
        case 25: *f1 = 5; *f2 = 5; break;
        case 10: *f1 = 2; *f2 = 5; break;
        case 20: *f1 = 4; *f2 = 5; break;
        case 32: *f1 = 4; *f2 = 8; break;
        default: *f1 = r; *f2 = 1; break;
        }
}

// stage tables of a plan over blocks of m points (FftPlan::tw_off): returns the entry count, fills `tab` when given
int build_stage_tables(FftPlan& pl, std::vector<float2>* tab)
{
    int off = 0, n_s = pl.m;
    for (int s = 0; s < pl.n_stages; s++)
        {
            int f1, f2;
            stage_factors(pl.radix[s], &f1, &f2);
            const int ms1 = n_s / f1;
            const bool prime = !is_small_radix(pl.radix[s]);
            const int jmax = prime ? n_s : (f2 > 1 ? ms1 / f2 : ms1);  // butterfly positions j < jmax (prime stage: exponents j2 * k < n_s)
            const int nh = (jmax + FFT_TW_STAGE_LO - 1) / FFT_TW_STAGE_LO;
            pl.tw_off[s] = off;
            pl.tw_nh[s] = nh;
            if (prime)
                {
                    // W_{n_s}^t two-level, then W_P^t for t < P
                    if (tab)
                        {
                            for (int l = 0; l < FFT_TW_STAGE_LO; l++)
                                {
                                    const double a = -TWO_PI * (double)l / (double)n_s;
                                    tab->push_back(make_float2((float)std::cos(a), (float)std::sin(a)));
                                }
                            for (int q = 0; q < nh; q++)
                                {
                                    const double a = -TWO_PI * (double)(q * FFT_TW_STAGE_LO) / (double)n_s;
                                    tab->push_back(make_float2((float)std::cos(a), (float)std::sin(a)));
                                }
                            for (int t = 0; t < pl.radix[s]; t++)
                                {
                                    const double a = -TWO_PI * (double)t / (double)pl.radix[s];
                                    tab->push_back(make_float2((float)std::cos(a), (float)std::sin(a)));
                                }
                        }
                    off += FFT_TW_STAGE_LO + nh + pl.radix[s];
                    n_s /= pl.radix[s];
                    continue;
                }
            const int n_tab = f2 > 1 ? 2 : 1;
            if (tab)
                for (int t = 0; t < n_tab; t++)
                    {
                        const double base = t == 0 ? (double)n_s : (double)ms1;  // W_{n_s}^j, W_{ms1}^j
                        for (int l = 0; l < FFT_TW_STAGE_LO; l++)
                            {
                                const double a = -TWO_PI * (double)l / base;
                                tab->push_back(make_float2((float)std::cos(a), (float)std::sin(a)));
                            }
                        for (int q = 0; q < nh; q++)
                            {
                                const double a = -TWO_PI * (double)(q * FFT_TW_STAGE_LO) / base;
                                tab->push_back(make_float2((float)std::cos(a), (float)std::sin(a)));
                            }
                    }
            off += n_tab * (FFT_TW_STAGE_LO + nh);
            n_s /= pl.radix[s];
        }
    pl.n_tw = off;
    return off;
}

// derived fields of a plan whose n, m, g_radix, n_stages and radix[] are set
void finish_plan(FftPlan& pl)
{
    const bool blk_map = getenv("GB200_ACQ_NO_BLKMAP") == nullptr;
    int n_s = pl.m;
    pl.map_blk = 0;
    for (int s = 0; s < pl.n_stages; s++)
        {
            const unsigned ms = (unsigned)(n_s / pl.radix[s]);  // butterfly positions per sub-block (ms2 of a fused stage)
            const unsigned nblk = (unsigned)(pl.m / n_s);
            pl.inv_ms[s] = ms == 1 ? 0u : (unsigned)((0x100000000ull + ms - 1) / ms);
            pl.inv_nblk[s] = nblk <= 1 ? 0u : (unsigned)((0x100000000ull + nblk - 1) / nblk);
            if (blk_map && ms > 1 && ms < 16 && (n_s & 1) && nblk >= 16) pl.map_blk |= 1u << s;
            n_s /= pl.radix[s];
        }
    pl.perm_rl = pl.radix[pl.n_stages - 1];
}

// threads per CTA: the largest stage radix decides how many butterflies a pass has
int plan_threads(const FftPlan& pl)
{
    int rmax = 2;
    for (int s = 0; s < pl.n_stages; s++) rmax = std::max(rmax, pl.radix[s]);
    const int bf = pl.m / std::min(rmax, 8);  // plain passes have m / R butterflies, fused ones m / (R1 R2)
    if (const char* e = getenv("GB200_ACQ_THREADS")) return std::min(ACQ_MAX_THREADS, std::max(64, atoi(e) / 32 * 32));  // tuning knob
    return std::min(ACQ_MAX_THREADS, std::max(128, (bf + 63) / 64 * 64));  // 512: the fused radix-25 pass holds 25 points per thread (128 registers)
}

int make_plan(gb200_acq* h, int max_smem)
{
    const int N = h->tlen;
    FftPlan& pl = h->plan;
    if (N < 8) return set_error(GB200_EINVAL, "fft_size %d: at least 8 points", N);
    pl.n = N;
    pl.n_hi = (N + FFT_TW_LO - 1) / FFT_TW_LO;
    const bool fuse = getenv("GB200_ACQ_PLAIN_RADIX") == nullptr;
    std::vector<int> radix;
    int g = 0;
    // smallest split that fits one SM's shared memory; a transform that fits only ONCE per SM is split in two when its halves fit twice
    // (two cells resident per SM, the halves of a cell joined through distributed shared memory: acq_inv_kernel<3>).
    // GB200_ACQ_SPLIT=<r> forces the split (tuning / tests).
    int r_first = 1;
    if (const char* e = getenv("GB200_ACQ_SPLIT")) r_first = std::max(1, std::min(8, atoi(e)));
    else if (N % 2 == 0 && N >= 8192 && getenv("GB200_ACQ_AUTOSPLIT") != nullptr)  // opt-in: measured slower on B200 (25000: 0.705 ms per C3 sweep against 0.459 unsplit; one more pass and the DSMEM stage outweigh the second resident cell)
        {
            std::vector<int> rx;
            FftPlan probe{};
            probe.m = N;
            if (factor_235(N, rx, fuse))
                {
                    probe.n_stages = (int)rx.size();
                    for (int s = 0; s < probe.n_stages; s++) probe.radix[s] = rx[s];
                    const size_t whole = ((size_t)N + build_stage_tables(probe, nullptr)) * sizeof(float2);
                    if (whole <= (size_t)max_smem && 2 * (whole + 1024) > (size_t)228 * 1024 && 2 * (whole / 2 + 4096) <= (size_t)228 * 1024) r_first = 2;
                }
        }
    // (the loop below runs once, or twice when its first answer is a split by 2 or 3 of a length that also splits by 4: see after it)
    const int r_first0 = r_first;
    for (int pass = 0; pass < 3; pass++)
    {
    for (int r = r_first; r <= 8 && !g; r++)
        {
            if (N % r) continue;
            pl.m = N / r;
            if (!factor_235(pl.m, radix, fuse)) return set_error(GB200_EINVAL, "fft_size %d: prime factors above 61 (or a single small factor next to them) are not supported", N);
            // the inverse kernel fuses the code product into its first stage and the search into its last: two stages at least
            if (radix.size() < 2) factor_235(pl.m, radix, false);
            if (radix.size() == 1 && radix[0] == 4) radix = {2, 2};
            if (radix.size() < 2) continue;
            pl.n_stages = (int)radix.size();
            for (int s = 0; s < pl.n_stages; s++) pl.radix[s] = radix[s];
            pl.y_len = 0;
            for (int s = 0; s < pl.n_stages; s++)
                if (!is_small_radix(pl.radix[s])) pl.y_len = pl.m;
            const size_t bytes = ((size_t)pl.m + pl.y_len + build_stage_tables(pl, nullptr)) * sizeof(float2);
            if (bytes <= (size_t)max_smem)
                {
                    g = r;
                    h->smem = bytes;
                }
        }
    // Split transforms below 65536 points: four blocks joined through distributed shared memory beat two (measured on the final tree, 32 PRN x
    // 80 bins, ms per sweep — 32000: 0.935 split 2 through global memory, 0.890 split 2 cluster, 0.835 split 4 cluster; 40000: 1.154 / 1.073 /
    // 0.995; 50000: 1.474 / 1.337 / 1.222; 64000: 2.204 / 1.945 / 1.946; split 5 wins at 32000 and 64000 but loses at 40000 and 50000, split 8 is
    // twice as slow): two or three CTAs per SM instead of one, and three-pass blocks.
    if (pass == 0 && g > 1 && g < 4 && N % 4 == 0 && N < 65536 && getenv("GB200_ACQ_SPLIT") == nullptr && getenv("GB200_ACQ_NO_SPLIT4") == nullptr)
        {
            g = 0;
            r_first = 4;
            continue;
        }
    if (pass == 1 && !g)  // (no split by 4 ... 8 fits after all: back to the first answer)
        {
            r_first = r_first0;
            continue;
        }
    break;
    }
    if (!g) return set_error(GB200_EINVAL, "fft_size %d: no split into <= 8 shared-memory blocks", N);
    pl.g_radix = g;
    pl.m = N / g;
    finish_plan(pl);
    h->tw_stage.clear();
    build_stage_tables(pl, &h->tw_stage);
    h->threads = plan_threads(pl);
    {
        // the kernels hold 128 registers per thread (512 threads fill an SM's register file): when several blocks fit an SM's
        // shared memory, split those 512 threads over 2 or 4 resident CTAs so that their load / barrier phases overlap
        // (measured: N = 4000 0.147 -> 0.095 ms per C1 sweep with 128 threads, N = 8000 0.192 -> 0.153 with 256)
        const size_t per_sm = 228 * 1024;
        const int fit = (int)(per_sm / (h->smem + 1024));
        const int k = fit >= 4 ? 4 : fit >= 2 ? 2 : 1;
        if (getenv("GB200_ACQ_THREADS") == nullptr) h->threads = k == 1 ? h->threads : std::max(128, std::min(h->threads, 512 / k));
    }
    // forward split: peel the first plain factor of the first stage off into the global stage
    h->split_fwd = false;
    if (g == 1 && N >= 8192 && getenv("GB200_ACQ_SPLIT_FWD") != nullptr)  // opt-in: measured 3 % slower than one CTA per bin on the C3 sweep
        {
            FftPlan& pf = h->plan_f;
            pf = pl;
            int f1, f2;
            stage_factors(pl.radix[0], &f1, &f2);
            pf.g_radix = f1;
            pf.m = N / f1;
            pf.n_stages = 0;
            if (f2 > 1) pf.radix[pf.n_stages++] = f2;
            for (int s = 1; s < pl.n_stages; s++) pf.radix[pf.n_stages++] = pl.radix[s];
            finish_plan(pf);
            h->tw_stage_f.clear();
            build_stage_tables(pf, &h->tw_stage_f);
            pf.y_len = pl.y_len ? pf.m : 0;
            h->smem_f = ((size_t)pf.m + pf.y_len + pf.n_tw) * sizeof(float2);
            h->threads_f = std::min(256, plan_threads(pf));  // 128 registers per thread: two CTAs per SM
            h->split_fwd = true;
        }
    // natural index -> position after the DIF stages (global stage first)
    h->pos_of_k.resize(N);
    for (int k = 0; k < N; k++)
        {
            int p = 0, size = N, kk = k;
            auto digit = [&](int r) {
                const int m = size / r;
                p += (kk % r) * m;
                kk /= r;
                size = m;
            };
            if (g > 1) digit(g);
            for (int s = 0; s < pl.n_stages; s++)
                {
                    int f1, f2;
                    stage_factors(pl.radix[s], &f1, &f2);
                    digit(f1);
                    if (f2 > 1) digit(f2);
                }
            h->pos_of_k[k] = p;
        }
    return GB200_OK;
}

unsigned long long turns_fix(double turns)
{
    double f = turns - std::floor(turns);
    if (f >= 1.0) f = 0.0;
    return (unsigned long long)std::ldexp(f, 53) << 11;
}

template <typename K>
cudaError_t opt_in_smem(K kernel, size_t smem)
{
    if (smem > 48 * 1024) return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    return cudaSuccess;
}

// forward transform of `batch` blocks: raw IQ + wipe-off (src_iq) or cf32 buffers; result in dst (digit-reversed order)
template <int IQ>
cudaError_t forward_iq(gb200_acq* h, int batch, float2* dst)
{
    const bool split = h->split_fwd;
    const FftPlan& pl = split ? h->plan_f : h->plan;
    const float2* twT = split ? h->d_twT_f : h->d_twT;
    const int threads = split ? h->threads_f : h->threads;
    const size_t smem = split ? h->smem_f : h->smem;
    cudaError_t e;
    if (pl.g_radix == 1)
        {
            if ((e = opt_in_smem(acq_fwd_kernel<IQ, true>, smem)) != cudaSuccess) return e;
            acq_fwd_kernel<IQ, true><<<dim3(1, batch), threads, smem, h->stream>>>(pl, h->d_iq, h->cfg.consumed_samples, h->d_dphi, h->exact_wipeoff ? nullptr : h->d_phase, twT, dst, 0);
            count_launch();
            return cudaGetLastError();
        }
    acq_gstage_fwd_kernel<IQ, true><<<dim3((pl.m + 255) / 256, batch), 256, 0, h->stream>>>(pl, h->d_iq, h->cfg.consumed_samples, h->d_dphi, h->exact_wipeoff ? nullptr : h->d_phase, h->d_twA, h->d_twB, h->d_stage);
    count_launch();
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if ((e = opt_in_smem(acq_fwd_kernel<0, false>, smem)) != cudaSuccess) return e;
    acq_fwd_kernel<0, false><<<dim3(pl.g_radix, batch), threads, smem, h->stream>>>(pl, h->d_stage, 0, nullptr, nullptr, twT, dst, 0);
    count_launch();
    return cudaGetLastError();
}

// forward transform of ONE cf32 block held in d_stage[0..N) (code replica / test hook); result in dst
cudaError_t forward_cf32(gb200_acq* h, float2* dst, int conj_out)
{
    const FftPlan& pl = h->plan;
    cudaError_t e;
    const float2* src = h->d_stage;
    if (pl.g_radix > 1)
        {
            // the global stage is not in place: stage into the second half of d_stage (it holds >= 2 blocks, see create)
            float2* tmp = h->d_stage + pl.n;
            acq_gstage_fwd_kernel<0, false><<<dim3((pl.m + 255) / 256, 1), 256, 0, h->stream>>>(pl, h->d_stage, 0, nullptr, nullptr, h->d_twA, h->d_twB, tmp);
            count_launch();
            if ((e = cudaGetLastError()) != cudaSuccess) return e;
            src = tmp;
        }
    if ((e = opt_in_smem(acq_fwd_kernel<0, false>, h->smem)) != cudaSuccess) return e;
    acq_fwd_kernel<0, false><<<dim3(pl.g_radix, 1), h->threads, h->smem, h->stream>>>(pl, src, 0, nullptr, nullptr, h->d_twT, dst, conj_out);
    count_launch();
    return cudaGetLastError();
}
}  // namespace

#include "acq_api.inc"
