// Hand-written batched mixed-radix FFT building blocks for the PCPS acquisition kernels (no cuFFT).
//
// Replaces FFTW reached through gr::fft::fft_complex_fwd/rev (src/algorithms/libs/gnss_sdr_fft.h:27-61; call sites
// pcps_acquisition.cc:113-114,206,647,653): unnormalised complex DFT of d_fft_size points, sizes such as 4000, 16000,
// 20000, 25000, 32000, 100000 (factors 2, 3, 5, 7).
//
// Scheme: in-place decimation-in-frequency forward transform whose output is left in mixed-radix digit-reversed order,
// element-wise product with a code spectrum stored in the SAME order, then the transposed decimation-in-time inverse
// that consumes that order and delivers natural order.  No reordering pass, no ping-pong buffer: a 25000-point block
// (200 KB) lives in one SM's shared memory.  Transforms longer than one SM's shared memory take their first DIF stage
// (last DIT stage) through global memory and continue as g_radix independent shared-memory blocks.
#pragma once
#include "common.cuh"

namespace gb200
{
#ifndef ACQ_MAX_THREADS
#define ACQ_MAX_THREADS 512  // threads per CTA of the shared-memory FFT kernels: 128 registers each fill the register file
#endif
constexpr int FFT_MAX_STAGES = 20;
constexpr int FFT_TW_SHIFT = 8;  // W_N^e = A[e >> 8] * B[e & 255]
constexpr int FFT_TW_LO = 1 << FFT_TW_SHIFT;

struct FftPlan
{
    int n;         // transform length N
    int m;         // shared-memory block length (N / g_radix)
    int g_radix;   // 1, or the radix of the stage done through global memory
    int n_stages;  // shared-memory stages
    int n_hi;      // entries of twiddle table A
    int radix[FFT_MAX_STAGES];      // DIF order
    unsigned inv_ms[FFT_MAX_STAGES];  // ceil(2^32 / ms) (0 when ms == 1) for the per-stage butterfly -> (block, j2) split
    // per-stage twiddle tables (shared-memory stages): stage s owns entries [tw_off[s], tw_off[s+1]) of the stage-table block,
    // laid out  lo1[32] hi1[tw_nh] (base W_{n_s})  and, for fused stages,  lo2[32] hi2[tw_nh] (base W_{n_s / R1}):
    // W^j = hi[j >> 5] * lo[j & 31], indexed by the butterfly position j, so a warp reads consecutive lo entries (no bank
    // conflicts) and one or two hi entries (broadcast)
    int tw_off[FFT_MAX_STAGES];
    int tw_nh[FFT_MAX_STAGES];
    // butterfly -> thread mapping: bit s set = consecutive threads take the same position j in consecutive sub-blocks (stride n_s,
    // odd: conflict-free, twiddles broadcast) instead of consecutive j; used when a sub-block has fewer than 16 positions
    unsigned map_blk;
    unsigned inv_nblk[FFT_MAX_STAGES];  // ceil(2^32 / (m / n_s)) for that mapping
    int n_tw;      // entries of the stage-table block
    int perm_rl;   // full radix of the last DIF stage: code spectra are stored [q][butterfly] within a block (see acq_inv_kernel)
    int y_len;     // m when the plan has a prime-radix stage (11 <= P <= 61: out-of-place, needs a second block in shared memory), else 0
};

constexpr int FFT_TW_STAGE_LO = 32;

__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ float2 cmulc(float2 a, float2 b) { return make_float2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y); }  // a * conj(b)
// complex add / subtract / scale as ONE packed f32x2 instruction each (FADD2 / FMUL2 / FFMA2; every lane rounds like the scalar op)
__device__ __forceinline__ float2 f2(f32x2 v)
{
    float2 r;
    up2(v, r.x, r.y);
    return r;
}
__device__ __forceinline__ f32x2 p2(float2 a) { return pk2(a.x, a.y); }
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return f2(add2(p2(a), p2(b))); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return f2(sub2(p2(a), p2(b))); }
__device__ __forceinline__ float2 cscale(float s, float2 a) { return f2(mul2(pk2(s, s), p2(a))); }                 // s * a
__device__ __forceinline__ float2 caxpy(float s, float2 a, float2 y) { return f2(fma2(pk2(s, s), p2(a), p2(y))); }  // s * a + y
// multiply by -i (forward) / +i (inverse)
template <bool INV>
__device__ __forceinline__ float2 rot90(float2 a)
{
    return INV ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x);
}

// W_N^e = exp(-2*pi*i*e/N), 0 <= e < N, from the two-level table staged in shared memory
__device__ __forceinline__ float2 fft_twiddle(const float2* sA, const float2* sB, int e)
{
    return cmul(sA[e >> FFT_TW_SHIFT], sB[e & (FFT_TW_LO - 1)]);
}

// W^j from one stage table (see FftPlan::tw_off)
__device__ __forceinline__ float2 fft_stage_twiddle(const float2* t, int j)
{
    return cmul(t[FFT_TW_STAGE_LO + (j >> 5)], t[j & (FFT_TW_STAGE_LO - 1)]);
}

// W_M^k = exp(-2*pi*i*k/M) as literals (k is a compile-time constant after unrolling)
template <int M>
__device__ __forceinline__ float2 fft_wconst(int k);
template <>
__device__ __forceinline__ float2 fft_wconst<7>(int k)
{
    switch (k)
        {
        case 1: return make_float2(0.62348980185873359f, -0.7818314824680298f);
        case 2: return make_float2(-0.22252093395631434f, -0.97492791218182362f);
        case 3: return make_float2(-0.90096886790241903f, -0.43388373911755823f);
        case 4: return make_float2(-0.90096886790241915f, 0.43388373911755801f);
        case 5: return make_float2(-0.22252093395631459f, 0.97492791218182362f);
        case 6: return make_float2(0.62348980185873337f, 0.78183148246802991f);
        default: return make_float2(1.0f, 0.0f);
        }
}
template <>
__device__ __forceinline__ float2 fft_wconst<8>(int k)
{
    return k == 1 ? make_float2(0.70710678118654757f, -0.70710678118654746f) : make_float2(1.0f, 0.0f);
}
template <>
__device__ __forceinline__ float2 fft_wconst<9>(int k)
{
    switch (k)
        {
        case 1: return make_float2(0.76604444311897801f, -0.64278760968653925f);
        case 2: return make_float2(0.17364817766693041f, -0.98480775301220802f);
        default: return make_float2(1.0f, 0.0f);
        }
}
// (2 x 5 and 4 x 5: a stray factor 2 or 4 fused with a factor 5 saves a whole pass for lengths such as 4000 = 16 * 10 * 25 and 8000 = 16 * 20 * 25)
template <>
__device__ __forceinline__ float2 fft_wconst<10>(int k)
{
    switch (k)
        {
        case 1: return make_float2(0.80901699437494745f, -0.58778525229247314f);
        case 2: return make_float2(0.30901699437494745f, -0.95105651629515353f);
        case 3: return make_float2(-0.30901699437494734f, -0.95105651629515364f);
        case 4: return make_float2(-0.80901699437494734f, -0.58778525229247325f);
        default: return make_float2(1.0f, 0.0f);
        }
}
template <>
__device__ __forceinline__ float2 fft_wconst<20>(int k)
{
    switch (k)
        {
        case 1: return make_float2(0.95105651629515353f, -0.3090169943749474f);
        case 2: return make_float2(0.80901699437494745f, -0.58778525229247314f);
        case 3: return make_float2(0.58778525229247314f, -0.80901699437494745f);
        case 4: return make_float2(0.30901699437494745f, -0.95105651629515353f);
        default: return make_float2(1.0f, 0.0f);
        }
}
// (4 x 8: 16000 = 32 * 20 * 25 and 20000 = 32 * 25 * 25 in three passes)
template <>
__device__ __forceinline__ float2 fft_wconst<32>(int k)
{
    switch (k)
        {
        case 1: return make_float2(0.98078528040323043f, -0.19509032201612825f);
        case 2: return make_float2(0.92387953251128674f, -0.38268343236508978f);
        case 3: return make_float2(0.83146961230254524f, -0.55557023301960218f);
        case 4: return make_float2(0.70710678118654757f, -0.70710678118654746f);
        case 5: return make_float2(0.55557023301960229f, -0.83146961230254524f);
        case 6: return make_float2(0.38268343236508984f, -0.92387953251128674f);
        case 7: return make_float2(0.19509032201612833f, -0.98078528040323043f);
        default: return make_float2(1.0f, 0.0f);
        }
}
template <>
__device__ __forceinline__ float2 fft_wconst<16>(int k)
{
    switch (k)
        {
        case 1: return make_float2(0.92387953251128674f, -0.38268343236508978f);
        case 2: return make_float2(0.70710678118654757f, -0.70710678118654746f);
        case 3: return make_float2(0.38268343236508984f, -0.92387953251128674f);
        default: return make_float2(1.0f, 0.0f);
        }
}
template <>
__device__ __forceinline__ float2 fft_wconst<25>(int k)
{
    switch (k)
        {
        case 1: return make_float2(0.96858316112863108f, -0.24868988716485479f);
        case 2: return make_float2(0.87630668004386358f, -0.48175367410171532f);
        case 3: return make_float2(0.72896862742141155f, -0.68454710592868862f);
        case 4: return make_float2(0.53582679497899655f, -0.84432792550201508f);
        default: return make_float2(1.0f, 0.0f);
        }
}

template <int R, bool INV>
struct Dft;

template <bool INV>
struct Dft<2, INV>
{
    static __device__ __forceinline__ void run(float2* v)
    {
        const float2 a = v[0], b = v[1];
        v[0] = cadd(a, b);
        v[1] = csub(a, b);
    }
};

template <bool INV>
struct Dft<3, INV>
{
    static __device__ __forceinline__ void run(float2* v)
    {
        const float2 t = cadd(v[1], v[2]);
        const float2 d = csub(v[1], v[2]);
        const float2 a = caxpy(-0.5f, t, v[0]);
        const float2 b = rot90<INV>(cscale(0.86602540378443865f, d));
        v[0] = cadd(v[0], t);
        v[1] = cadd(a, b);
        v[2] = csub(a, b);
    }
};

template <bool INV>
struct Dft<4, INV>
{
    static __device__ __forceinline__ void run(float2* v)
    {
        const float2 t0 = cadd(v[0], v[2]), t1 = csub(v[0], v[2]);
        const float2 t2 = cadd(v[1], v[3]), t3 = rot90<INV>(csub(v[1], v[3]));
        v[0] = cadd(t0, t2);
        v[2] = csub(t0, t2);
        v[1] = cadd(t1, t3);
        v[3] = csub(t1, t3);
    }
};

// 8 points: one radix-2 level with the W_8 rotations, then two 4-point butterflies (even outputs from the sums, odd from the differences)
template <bool INV>
struct Dft<8, INV>
{
    static __device__ __forceinline__ void run(float2* v)
    {
        const float h = 0.70710678118654757f;
        float2 a[4], b[4];
#pragma unroll
        for (int k = 0; k < 4; k++)
            {
                a[k] = cadd(v[k], v[k + 4]);
                b[k] = csub(v[k], v[k + 4]);
            }
        // b1 * W_8, b2 * W_8^2, b3 * W_8^3 (conjugates for the inverse)
        b[1] = INV ? make_float2(h * (b[1].x - b[1].y), h * (b[1].x + b[1].y)) : make_float2(h * (b[1].x + b[1].y), h * (b[1].y - b[1].x));
        b[2] = rot90<INV>(b[2]);
        b[3] = INV ? make_float2(h * (-b[3].x - b[3].y), h * (b[3].x - b[3].y)) : make_float2(h * (b[3].y - b[3].x), h * (-b[3].x - b[3].y));
        Dft<4, INV>::run(a);
        Dft<4, INV>::run(b);
#pragma unroll
        for (int k = 0; k < 4; k++)
            {
                v[2 * k] = a[k];
                v[2 * k + 1] = b[k];
            }
    }
};

template <bool INV>
struct Dft<5, INV>
{
    static __device__ __forceinline__ void run(float2* v)
    {
        const float c1 = 0.30901699437494742f, c2 = -0.80901699437494742f;
        const float s1 = 0.95105651629515357f, s2 = 0.58778525229247313f;
        const float2 t1 = cadd(v[1], v[4]), t2 = cadd(v[2], v[3]);
        const float2 t3 = csub(v[1], v[4]), t4 = csub(v[2], v[3]);
        const float2 a1 = caxpy(c2, t2, caxpy(c1, t1, v[0]));
        const float2 a2 = caxpy(c1, t2, caxpy(c2, t1, v[0]));
        const float2 b1 = rot90<INV>(caxpy(s2, t4, cscale(s1, t3)));
        const float2 b2 = rot90<INV>(caxpy(-s1, t4, cscale(s2, t3)));
        v[0] = cadd(cadd(v[0], t1), t2);
        v[1] = cadd(a1, b1);
        v[4] = csub(a1, b1);
        v[2] = cadd(a2, b2);
        v[3] = csub(a2, b2);
    }
};

// O(R^2) radix-7 butterfly with literal coefficients W_7^q
template <bool INV>
struct Dft<7, INV>
{
    static __device__ __forceinline__ void run(float2* v)
    {
        float2 y[7];
#pragma unroll
        for (int k = 0; k < 7; k++)
            {
                float2 acc = v[0];
#pragma unroll
                for (int j = 1; j < 7; j++)
                    {
                        const float2 w = fft_wconst<7>((j * k) % 7);
                        acc = cadd(acc, INV ? cmulc(v[j], w) : cmul(v[j], w));
                    }
                y[k] = acc;
            }
#pragma unroll
        for (int k = 0; k < 7; k++) v[k] = y[k];
    }
};

// Hooks of a stage.  Pre: multiplier applied to the points as they are read (only on the first inverse stage, whose butterflies
// own R contiguous points: point q of butterfly id).  Post: consumer that replaces the store (only on the last inverse stage,
// whose outputs are in natural order: y[n]).
struct FftNoPre
{
    static constexpr bool on = false;
    __device__ __forceinline__ float2 operator()(int, int) const { return make_float2(1.0f, 0.0f); }
};
struct FftNoPost
{
    static constexpr bool batched = false;
    static constexpr bool on = false;
    __device__ __forceinline__ void operator()(float2, int) {}
    template <int R>
    __device__ __forceinline__ void batch(const float2*, int, int)
    {
    }
};

// One in-place stage over a shared-memory block x[0..m_blk): sub-blocks of n_s points, radix R; T = the stage's table.
// Forward (DIF): butterfly, then twiddle W_{n_s}^{j2*k}.  Inverse (DIT): conj twiddle, then inverse butterfly.
template <int R, bool INV, class Pre, class Post>
__device__ __forceinline__ void fft_smem_stage(float2* x, int m_blk, int n_s, unsigned inv_ms, unsigned inv_nblk, const float2* T, const Pre& pre, Post& post)
{
    const int ms = n_s / R;
    const int nb = m_blk / R;
    const int nblk = m_blk / n_s;
#pragma unroll(Pre::on ? 2 : 1)
    for (int id = threadIdx.x; id < nb; id += blockDim.x)
        {
            int blk, j2;
            if (inv_nblk)
                {
                    j2 = (int)__umulhi((unsigned)id, inv_nblk);  // id / nblk
                    blk = id - j2 * nblk;
                }
            else
                {
                    blk = ms == 1 ? id : (int)__umulhi((unsigned)id, inv_ms);  // id / ms (exact for id * ms < 2^32)
                    j2 = id - blk * ms;
                }
            float2* p = x + blk * n_s + j2;
            float2 v[R];
#pragma unroll
            for (int k = 0; k < R; k++) v[k] = Pre::on ? cmul(p[k * ms], pre(k, id)) : p[k * ms];
            if (INV && ms > 1)
                {
                    const float2 w1 = fft_stage_twiddle(T, j2);
                    float2 w = w1;
#pragma unroll
                    for (int k = 1; k < R; k++)
                        {
                            v[k] = cmulc(v[k], w);
                            if (k + 1 < R) w = cmul(w, w1);
                        }
                }
            Dft<R, INV>::run(v);
            if (!INV && ms > 1)
                {
                    const float2 w1 = fft_stage_twiddle(T, j2);
                    float2 w = w1;
#pragma unroll
                    for (int k = 1; k < R; k++)
                        {
                            v[k] = cmul(v[k], w);
                            if (k + 1 < R) w = cmul(w, w1);
                        }
                }
            if (Post::on && Post::batched)
                post.template batch<R>(v, blk * n_s + j2, ms);  // the butterfly's R outputs: indices first + k * ms, ascending
            else
                {
#pragma unroll
                    for (int k = 0; k < R; k++)
                        {
                            if (Post::on)
                                post(v[k], blk * n_s + j2 + k * ms);
                            else
                                p[k * ms] = v[k];
                        }
                }
        }
    if (!Post::on) __syncthreads();
}

// Two consecutive in-place stages (radix R1 on sub-blocks of n_s points, then radix R2 on their n_s / R1 point parts)
// fused in registers: one pass over shared memory and one barrier instead of two.  The data layout before and after is
// exactly that of the two separate stages, so plans may mix fused and plain stages freely.  A thread owns the R1 * R2
// points  base + k * ms1 + k2 * ms2  (k < R1, k2 < R2).  First-stage twiddles of column k2 are
// W_{n_s}^{(k2*ms2 + j) k} = (W_{n_s}^j * W_{R1 R2}^{k2})^k : one table look-up and R2 - 1 literal rotations per butterfly.
template <int R1, int R2, bool INV, class Pre, class Post>
__device__ __forceinline__ void fft_smem_stage2(float2* x, int m_blk, int n_s, unsigned inv_ms, unsigned inv_nblk, const float2* T, int nh, const Pre& pre, Post& post)
{
    const int ms1 = n_s / R1, ms2 = ms1 / R2;
    const int nb = m_blk / (R1 * R2);
    const int nblk = m_blk / n_s;
    const float2* T2 = T + FFT_TW_STAGE_LO + nh;
    for (int id = threadIdx.x; id < nb; id += blockDim.x)
        {
            int blk, j;
            if (inv_nblk)
                {
                    j = (int)__umulhi((unsigned)id, inv_nblk);  // id / nblk
                    blk = id - j * nblk;
                }
            else
                {
                    blk = ms2 == 1 ? id : (int)__umulhi((unsigned)id, inv_ms);  // id / ms2
                    j = id - blk * ms2;
                }
            float2* p = x + blk * n_s + j;
            float2 v[R1][R2];
#pragma unroll
            for (int k = 0; k < R1; k++)
#pragma unroll
                for (int k2 = 0; k2 < R2; k2++) v[k][k2] = Pre::on ? cmul(p[k * ms1 + k2 * ms2], pre(k * R2 + k2, id)) : p[k * ms1 + k2 * ms2];
            // second-stage twiddles W_{ms1}^{j * k2} (shared by all k)
            float2 wb1 = make_float2(1.0f, 0.0f);
            float2 w10 = make_float2(1.0f, 0.0f);
            if (ms2 > 1)
                {
                    w10 = fft_stage_twiddle(T, j);
                    wb1 = fft_stage_twiddle(T2, j);
                }
            if (INV)
                {
                    if (ms2 > 1)
                        {
                            float2 w = wb1;
#pragma unroll
                            for (int k2 = 1; k2 < R2; k2++)
                                {
#pragma unroll
                                    for (int k = 0; k < R1; k++) v[k][k2] = cmulc(v[k][k2], w);
                                    if (k2 + 1 < R2) w = cmul(w, wb1);
                                }
                        }
#pragma unroll
                    for (int k = 0; k < R1; k++) Dft<R2, true>::run(v[k]);
                }
#pragma unroll
            for (int k2 = 0; k2 < R2; k2++)
                {
                    const float2 wc = fft_wconst<R1 * R2>(k2);
                    const float2 w1 = k2 == 0 ? w10 : (ms2 > 1 ? cmul(w10, wc) : wc);
                    float2 t[R1];
#pragma unroll
                    for (int k = 0; k < R1; k++) t[k] = v[k][k2];
                    if (INV && (ms2 > 1 || k2 > 0))
                        {
                            float2 w = w1;
#pragma unroll
                            for (int k = 1; k < R1; k++)
                                {
                                    t[k] = cmulc(t[k], w);
                                    if (k + 1 < R1) w = cmul(w, w1);
                                }
                        }
                    Dft<R1, INV>::run(t);
                    if (!INV && (ms2 > 1 || k2 > 0))
                        {
                            float2 w = w1;
#pragma unroll
                            for (int k = 1; k < R1; k++)
                                {
                                    t[k] = cmul(t[k], w);
                                    if (k + 1 < R1) w = cmul(w, w1);
                                }
                        }
#pragma unroll
                    for (int k = 0; k < R1; k++) v[k][k2] = t[k];
                }
            if (!INV)
                {
#pragma unroll
                    for (int k = 0; k < R1; k++) Dft<R2, false>::run(v[k]);
                    if (ms2 > 1)
                        {
                            float2 w = wb1;
#pragma unroll
                            for (int k2 = 1; k2 < R2; k2++)
                                {
#pragma unroll
                                    for (int k = 0; k < R1; k++) v[k][k2] = cmul(v[k][k2], w);
                                    if (k2 + 1 < R2) w = cmul(w, wb1);
                                }
                        }
                }
            if (Post::on && Post::batched)
                post.template batch<R1 * R2>(&v[0][0], blk * n_s + j, ms2);  // k * ms1 + k2 * ms2 = (k * R2 + k2) * ms2: ascending in storage order
            else
                {
#pragma unroll
                    for (int k = 0; k < R1; k++)
#pragma unroll
                        for (int k2 = 0; k2 < R2; k2++)
                            {
                                if (Post::on)
                                    post(v[k][k2], blk * n_s + j + k * ms1 + k2 * ms2);
                                else
                                    p[k * ms1 + k2 * ms2] = v[k][k2];
                            }
                }
        }
    if (!Post::on) __syncthreads();
}

// Stage of prime radix P (11 ... 61), for transform lengths such as 6625 = 5^3 * 53 (6.625 MS/s front-ends) or
// 5456 = 2^4 * 11 * 31: the DFT by definition, one thread per OUTPUT point, out of place through y (the second block in
// shared memory), O(P) per point.  T: two-level table of W_{n_s}^t for t < n_s, followed by W_P^t, t < P.  Never the first
// or the last stage of a plan (the hooks live in the small-radix stages), so it takes none.
template <bool INV>
__device__ __noinline__ void fft_smem_stage_prime(float2* x, float2* y, int m_blk, int n_s, int P, const float2* T, int nh)
{
    const int ms = n_s / P;
    const float2* Cf = T + FFT_TW_STAGE_LO + nh;  // W_P^t
    if (INV)
        {
            // DIT: conj twiddle W_{n_s}^{j2 k} on the inputs first
            for (int id = threadIdx.x; id < m_blk; id += blockDim.x)
                {
                    const int r = id % n_s, k = r / ms, j2 = r - k * ms;
                    if (k > 0 && j2 > 0) x[id] = cmulc(x[id], fft_stage_twiddle(T, j2 * k));
                }
            __syncthreads();
        }
    for (int id = threadIdx.x; id < m_blk; id += blockDim.x)
        {
            const int blk = id / n_s, r = id - blk * n_s, q = r / ms, j2 = r - q * ms;
            const float2* p = x + blk * n_s + j2;
            float2 acc = p[0];
            int e = 0;  // (q * k) mod P
            for (int k = 1; k < P; k++)
                {
                    e += q;
                    if (e >= P) e -= P;
                    const float2 c = Cf[e];
                    const float2 v = p[k * ms];
                    acc = cadd(acc, INV ? cmulc(v, c) : cmul(v, c));
                }
            y[id] = acc;
        }
    __syncthreads();
    for (int id = threadIdx.x; id < m_blk; id += blockDim.x)
        {
            float2 v = y[id];
            if (!INV)
                {
                    const int r = id % n_s, q = r / ms, j2 = r - q * ms;
                    if (q > 0 && j2 > 0) v = cmul(v, fft_stage_twiddle(T, j2 * q));  // DIF: twiddle W_{n_s}^{j2 q} on the outputs
                }
            x[id] = v;
        }
    __syncthreads();
}

template <bool INV, class Pre, class Post>
__device__ __forceinline__ void fft_smem_stage_any(const FftPlan& pl, int s, float2* x, int n_s, const float2* sT, const Pre& pre, Post& post)
{
    const unsigned inv_ms = pl.inv_ms[s], inv_nblk = ((pl.map_blk >> s) & 1u) ? pl.inv_nblk[s] : 0u;
    const float2* T = sT + pl.tw_off[s];
    const int nh = pl.tw_nh[s], m_blk = pl.m;
    switch (pl.radix[s])
        {
        case 2: fft_smem_stage<2, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, pre, post); break;
        case 3: fft_smem_stage<3, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, pre, post); break;
        case 4: fft_smem_stage<4, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, pre, post); break;
        case 5: fft_smem_stage<5, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, pre, post); break;
        case 8: fft_smem_stage2<4, 2, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, nh, pre, post); break;  // fused pairs: same layout as the two plain stages
        case 9: fft_smem_stage2<3, 3, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, nh, pre, post); break;
        case 16: fft_smem_stage2<4, 4, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, nh, pre, post); break;
        case 25: fft_smem_stage2<5, 5, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, nh, pre, post); break;
        case 10: fft_smem_stage2<2, 5, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, nh, pre, post); break;
        case 20: fft_smem_stage2<4, 5, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, nh, pre, post); break;
        case 32: fft_smem_stage2<4, 8, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, nh, pre, post); break;
        case 7: fft_smem_stage<7, INV>(x, m_blk, n_s, inv_ms, inv_nblk, T, pre, post); break;
        default: fft_smem_stage_prime<INV>(x, x + pl.m, m_blk, n_s, pl.radix[s], T, nh); break;  // prime radix 11 ... 61 (never with hooks)
        }
}

// all shared-memory stages of one block of plan.m points, forward: natural -> digit-reversed
__device__ __forceinline__ void fft_smem_forward(const FftPlan& pl, float2* x, const float2* sT)
{
    int n_s = pl.m;
    FftNoPre pre;
    FftNoPost post;
    for (int s = 0; s < pl.n_stages; s++)
        {
            fft_smem_stage_any<false>(pl, s, x, n_s, sT, pre, post);
            n_s /= pl.radix[s];
        }
}

// inverse: digit-reversed -> natural (stages in reverse order).  `pre` multiplies the input inside the first stage and
// `post` consumes the natural-order output of the last one; with a single stage both hooks cannot be fused at once, so
// the caller checks fft_can_fuse() first.
__device__ __forceinline__ bool fft_can_fuse(const FftPlan& pl) { return pl.n_stages >= 2; }

template <class Pre, class Post>
__device__ __forceinline__ void fft_smem_inverse(const FftPlan& pl, float2* x, const float2* sT, const Pre& pre, Post& post)
{
    FftNoPre no_pre;
    FftNoPost no_post;
    int n_s = 1;
    for (int s = pl.n_stages - 1; s >= 0; s--)
        {
            n_s *= pl.radix[s];
            if (Pre::on && s == pl.n_stages - 1)
                fft_smem_stage_any<true>(pl, s, x, n_s, sT, pre, no_post);
            else if (Post::on && s == 0)
                fft_smem_stage_any<true>(pl, s, x, n_s, sT, no_pre, post);
            else
                fft_smem_stage_any<true>(pl, s, x, n_s, sT, no_pre, no_post);
        }
}

// stage tables global -> shared (all threads), followed by a barrier in the caller
__device__ __forceinline__ void fft_load_tables(const FftPlan& pl, const float2* gT, float2* sT)
{
    for (int i = threadIdx.x; i < pl.n_tw; i += blockDim.x) sT[i] = gT[i];
}
}  // namespace gb200
