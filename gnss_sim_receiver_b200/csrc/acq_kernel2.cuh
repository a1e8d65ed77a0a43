// Inverse side of the acquisition search (included by acq_kernel.cuh).
#pragma once
#include <cooperative_groups.h>

namespace gb200
{
// grid = (g_radix, n_bins, n_slots).  One CTA transforms block blockIdx.x of  X[bin] .* C[slot]  back to code phase:
//   X[bin] is staged in shared memory by one TMA bulk copy (no register staging) while the threads load the stage tables;
//   the product with the code spectrum happens inside the first inverse stage (coalesced loads of the permuted spectrum);
//   MODE 0: (g_radix == 1) the last inverse stage feeds |.|^2, grid accumulation, max / argmax / sum straight from
//           registers                                                      -> cells[(slot_i * n_bins + bin)]
//   MODE 1: (g_radix  > 1) store the block to scratch[(slot_i * n_bins + bin)][N] for the global inverse stage
//   MODE 2: test hook — no code product, store the natural-order complex output to scratch
//   MODE 3: (g_radix  > 1) the g_radix CTAs of a cell form a THREAD-BLOCK CLUSTER: each transforms its block in its own shared memory,
//           then the last (radix g_radix) stage of the transform is taken straight out of the partners' shared memory (DSMEM): CTA r
//           combines the butterflies j2 in its slice of [0, m), feeds |.|^2 / grid / max / sum and writes part r of the cell.  Nothing
//           is staged in global memory; with two blocks of 12500 points, two cells are resident per SM for N = 25000.
template <int MODE>
__global__ void __launch_bounds__(ACQ_MAX_THREADS) acq_inv_kernel(AcqParams p, const float2* __restrict__ X, const float2* __restrict__ Cs, const int* __restrict__ slots,
    const float2* __restrict__ gT, float2* __restrict__ scratch, float* __restrict__ grid, AcqCell* __restrict__ cells, const float2* __restrict__ gA, const float2* __restrict__ gB)
{
    extern __shared__ __align__(16) float2 sm[];
    const FftPlan& pl = p.plan;
    float2* x = sm;
    float2* sT = sm + pl.m + pl.y_len;
    const int bin = blockIdx.y, si = blockIdx.z;
    const size_t blk_off = (size_t)blockIdx.x * pl.m;
    const float2* xs = X + (size_t)bin * pl.n + blk_off;
    __shared__ __align__(8) uint64_t s_bar;
    const bool bulk = (pl.m & 1) == 0 && ((size_t)pl.n & 1) == 0;  // 16-byte alignment of every block start and length
    if (bulk)
        {
            if (threadIdx.x == 0)
                {
                    mbar_init(&s_bar, 1);
                    fence_barrier_init();
                    mbar_expect_tx(&s_bar, (uint32_t)pl.m * 8u);
                    bulk_g2s(x, xs, (uint32_t)pl.m * 8u, &s_bar);
                }
        }
    else
        {
            for (int i = threadIdx.x; i < pl.m; i += blockDim.x) x[i] = xs[i];
        }
    fft_load_tables(pl, gT, sT);
    __syncthreads();  // tables (and the plain copy) visible; the barrier word is initialised before anyone polls it
    if (bulk) mbar_wait(&s_bar, 0);
    const size_t cell = (size_t)si * p.n_bins + bin;
    if (MODE == 2)
        {
            FftNoPre pre;
            FftNoPost post;
            fft_smem_inverse(pl, x, sT, pre, post);
        }
    else
        {
            AcqCodeProduct pre;
            pre.c = Cs + (size_t)slots[si] * pl.n + blk_off;
            pre.nb = pl.m / pl.perm_rl;
            if (MODE == 1 || MODE == 3)
                {
                    FftNoPost post;
                    fft_smem_inverse(pl, x, sT, pre, post);
                }
            else
                {
                    AcqSink sink;
                    sink.out_offset = p.out_offset;
                    sink.neff = p.neff;
                    sink.accumulate = p.dwell_index > 0;
                    sink.grid_row = p.keep_grid ? grid + ((size_t)slots[si] * p.n_bins + bin) * p.neff : nullptr;
                    sink.fast = (!p.keep_grid && p.out_offset == 0 && p.neff == pl.n) ? 1 : 0;
                    sink.best = -1.0f;
                    sink.sum = 0.0f;
                    sink.besti = 0x7fffffff;
                    fft_smem_inverse(pl, x, sT, pre, sink);
                    acq_block_reduce(sink.best, sink.besti, sink.sum);
                    if (threadIdx.x == 0)
                        {
                            AcqCell c;
                            c.vmax = sink.best;
                            c.imax = sink.besti;
                            c.sum = sink.sum;
                            c.pad_ = 0;
                            cells[cell] = c;
                        }
                    return;
                }
        }
    if (MODE == 3)
        {
            namespace cg = cooperative_groups;
            cg::cluster_group cluster = cg::this_cluster();
            cluster.sync();  // every block of the cell is transformed
            const int R = pl.g_radix, r = (int)cluster.block_rank(), unit = pl.n / R;
            const float2* xb[8];
#pragma unroll
            for (int k1 = 0; k1 < 8; k1++) xb[k1] = k1 < R ? cluster.map_shared_rank(x, k1) : x;
            AcqSink sink;
            sink.out_offset = p.out_offset;
            sink.fast = (!p.keep_grid && p.out_offset == 0 && p.neff == pl.n) ? 1 : 0;
            sink.neff = p.neff;
            sink.accumulate = p.dwell_index > 0;
            sink.grid_row = p.keep_grid ? grid + ((size_t)slots[si] * p.n_bins + bin) * p.neff : nullptr;
            sink.best = -1.0f;
            sink.sum = 0.0f;
            sink.besti = 0x7fffffff;
            const int per = (pl.m + R - 1) / R;
            const int lo = r * per, hi = min(pl.m, lo + per);
#pragma unroll 2
            for (int j2 = lo + threadIdx.x; j2 < hi; j2 += blockDim.x)
                {
                    // y[j1 * m + j2] = sum_k1 conj(W_N^{j2 k1}) z_k1[j2] conj(W_R^{j1 k1})   (last DIT stage)
                    float2 v[8];
#pragma unroll
                    for (int k1 = 0; k1 < 8; k1++)
                        if (k1 < R)
                            {
                                const float2 t = xb[k1][j2];
                                v[k1] = k1 ? cmulc(t, fft_twiddle(gA, gB, (int)((long long)j2 * k1))) : t;
                            }
                    switch (R)
                        {
                        case 2: Dft<2, true>::run(v); break;
                        case 3: Dft<3, true>::run(v); break;
                        case 4: Dft<4, true>::run(v); break;
                        case 5: Dft<5, true>::run(v); break;
                        case 7: Dft<7, true>::run(v); break;
                        default:
                            {
                                float2 y[8];
                                for (int j1 = 0; j1 < R; j1++)
                                    {
                                        float2 acc = v[0];
                                        for (int k1 = 1; k1 < R; k1++) acc = cadd(acc, cmulc(v[k1], fft_twiddle(gA, gB, ((j1 * k1) % R) * unit)));
                                        y[j1] = acc;
                                    }
                                for (int j1 = 0; j1 < R; j1++) v[j1] = y[j1];
                            }
                        }
                    // outputs j1 * m + j2, ascending in j1: the per-butterfly form of the search (AcqSink::batch) for the radices a split uses
                    switch (R)
                        {
                        case 2: sink.template batch<2>(v, j2, pl.m); break;
                        case 4: sink.template batch<4>(v, j2, pl.m); break;
                        case 5: sink.template batch<5>(v, j2, pl.m); break;
                        case 8: sink.template batch<8>(v, j2, pl.m); break;
                        default:
#pragma unroll
                            for (int j1 = 0; j1 < 8; j1++)
                                if (j1 < R) sink(v[j1], j1 * pl.m + j2);
                        }
                }
            acq_block_reduce(sink.best, sink.besti, sink.sum);
            if (threadIdx.x == 0)
                {
                    AcqCell c;
                    c.vmax = sink.best;
                    c.imax = sink.besti;
                    c.sum = sink.sum;
                    c.pad_ = 0;
                    cells[cell * p.n_parts + r] = c;
                }
            cluster.sync();  // no block leaves while a partner may still read its shared memory
            return;
        }
    float2* d = scratch + cell * pl.n + blk_off;
    for (int i = threadIdx.x; i < pl.m; i += blockDim.x) d[i] = x[i];
}

// Last DIT stage through global memory: y[j1*m + j2] = sum_k1 conj(W_N^{j2 k1}) z_k1[j2] conj(W_R^{j1 k1}), followed by the
// same epilogue.  grid = (n_parts, n_bins, n_slots); part = a contiguous range of j2.
__global__ void __launch_bounds__(256) acq_gstage_inv_kernel(AcqParams p, const float2* __restrict__ scratch, const int* __restrict__ slots,
    const float2* __restrict__ gA, const float2* __restrict__ gB, float* __restrict__ grid, AcqCell* __restrict__ cells, float2* __restrict__ raw_out)
{
    const FftPlan& pl = p.plan;
    const int R = pl.g_radix, unit = pl.n / R;
    const int bin = blockIdx.y, si = blockIdx.z;
    const size_t cell = (size_t)si * p.n_bins + bin;
    const float2* z = scratch + cell * pl.n;
    float* grow = (p.keep_grid && !raw_out) ? grid + ((size_t)slots[si] * p.n_bins + bin) * p.neff : nullptr;
    const int per = (pl.m + gridDim.x - 1) / gridDim.x;
    const int lo = blockIdx.x * per, hi = min(pl.m, lo + per);
    float best = -1.0f, sum = 0.0f;
    int besti = 0x7fffffff;
    // a thread owns whole butterflies; the R outputs of one butterfly are R samples m apart, visited in increasing n
    // across the j1 loop only per thread, so ties are resolved by (value, index) in the reduction
    for (int j2 = lo + threadIdx.x; j2 < hi; j2 += blockDim.x)
        {
            float2 v[8];
            for (int k1 = 0; k1 < R; k1++)
                {
                    const float2 t = z[k1 * pl.m + j2];
                    v[k1] = k1 ? cmulc(t, fft_twiddle(gA, gB, (int)((long long)j2 * k1))) : t;
                }
            for (int j1 = 0; j1 < R; j1++)
                {
                    float2 acc = v[0];
                    for (int k1 = 1; k1 < R; k1++) acc = cadd(acc, cmulc(v[k1], fft_twiddle(gA, gB, ((j1 * k1) % R) * unit)));
                    const int n = j1 * pl.m + j2;
                    if (raw_out)
                        raw_out[cell * pl.n + n] = acc;
                    else
                        {
                            // strict '>' inside acq_consume keeps the first maximum only if visited in increasing index
                            // order; here indices of one thread are not monotone across j2, so compare (value, index)
                            const int i = n - p.out_offset;
                            if (i >= 0 && i < p.neff)
                                {
                                    float g = __fadd_rn(__fmul_rn(acc.x, acc.x), __fmul_rn(acc.y, acc.y));  // as AcqSink: no FMA contraction
                                    if (grow)
                                        {
                                            if (p.dwell_index > 0) g = __fadd_rn(grow[i], g);
                                            grow[i] = g;
                                        }
                                    sum += g;
                                    if (g > best || (g == best && i < besti))
                                        {
                                            best = g;
                                            besti = i;
                                        }
                                }
                        }
                }
        }
    if (raw_out) return;
    acq_block_reduce(best, besti, sum);
    if (threadIdx.x == 0)
        {
            AcqCell c;
            c.vmax = best;
            c.imax = besti;
            c.sum = sum;
            c.pad_ = 0;
            cells[cell * p.n_parts + blockIdx.x] = c;
        }
}

// One CTA per slot: maximum over (bin, part) with "first Doppler bin, then first code phase" tie rule
// (pcps_acquisition.cc:505-514), then the test statistic and the Gnss_Synchro fields (:518-527 / :565-596, :691-696).
__global__ void __launch_bounds__(256) acq_stats_kernel(AcqParams p, const AcqCell* __restrict__ cells, const int* __restrict__ slots, const float* __restrict__ grid,
    gb200_acq_result* __restrict__ results)
{
    __shared__ float s_best;
    __shared__ int s_key;
    const int si = blockIdx.x;
    const AcqCell* c = cells + (size_t)si * p.n_bins * p.n_parts;
    // key = bin * neff + index : lexicographic (bin, index) order
    float best = -1.0f, dummy = 0.0f;
    int key = 0x7fffffff;
    for (int i = threadIdx.x; i < p.n_bins * p.n_parts; i += blockDim.x)
        {
            const int bin = i / p.n_parts;
            const float v = c[i].vmax;
            const int k = bin * p.neff + c[i].imax;
            // the reference starts from grid_maximum = 0 and uses strict '>': an all-zero grid leaves (0, 0)
            if (v > best || (v == best && k < key))
                {
                    best = v;
                    key = k;
                }
        }
    acq_block_reduce(best, key, dummy);
    if (threadIdx.x == 0)
        {
            if (!(best > 0.0f))
                {
                    best = 0.0f;
                    key = 0;
                }
            s_best = best;
            s_key = key;
        }
    __syncthreads();
    best = s_best;
    key = s_key;
    const int bin = key / p.neff, idx = key - bin * p.neff;
    float denom = 0.0f;
    if (p.use_cfar)
        {
            const int opp = (bin + p.n_bins / 2) % p.n_bins;  // :518
            float s = 0.0f;
            if (threadIdx.x == 0)
                {
                    for (int k = 0; k < p.n_parts; k++) s += c[opp * p.n_parts + k].sum;
                    denom = (float)((double)(s / (float)p.neff) / 2.0 / p.num_dwells);  // :519 (float / int, then double)
                }
        }
    else
        {
            // second peak of the winning bin outside +-samples_per_chip around the first (:565-593); the reference scans
            // d_fft_size entries of which only the first neff are ever written, the rest are zero
            const int nfull = p.plan.n;
            int ex1 = idx - p.samples_per_chip, ex2 = idx + p.samples_per_chip;
            if (ex1 < 0)
                ex1 = nfull + ex1;
            else if (ex2 >= nfull)
                ex2 = ex2 - nfull;
            const float* row = grid + ((size_t)slots[si] * p.n_bins + bin) * p.neff;
            float b2 = -1.0f, d2 = 0.0f;
            int i2 = 0x7fffffff;
            for (int i = threadIdx.x; i < p.neff; i += blockDim.x)
                {
                    // excluded: the circular range [ex1, ex2) walked by the reference's do/while
                    const bool excl = ex1 <= ex2 ? (i >= ex1 && i < ex2) : (i >= ex1 || i < ex2);
                    const float v = excl ? 0.0f : row[i];
                    if (v > b2)
                        {
                            b2 = v;
                            i2 = i;
                        }
                }
            acq_block_reduce(b2, i2, d2);
            denom = b2;
        }
    if (threadIdx.x == 0)
        {
            gb200_acq_result r;
            r.indext = (unsigned)idx;
            r.doppler_index = bin;
            r.doppler_hz = p.doppler_base + p.doppler_step * bin;
            r.peak = best;
            r.input_power = denom;
            r.test_statistics = best / denom;
            r.acq_delay_samples = fmodf((float)idx, p.samples_per_code);
            r.acq_doppler_hz = (float)r.doppler_hz;
            results[si] = r;
        }
}
}  // namespace gb200
