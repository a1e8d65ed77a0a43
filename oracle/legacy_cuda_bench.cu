// TEST / BASELINE INFRASTRUCTURE — not part of the product.
//
// Driver for the ONE GPU kernel the reference ships for the tracking hot path: cuda_multicorrelator
// (src/algorithms/tracking/libs/cuda_multicorrelator.{h,cu}), compiled UNMODIFIED from /root/reference where it lies
// (oracle/Makefile target _ref/legacy_cuda_bench; only the arch flag of src/algorithms/tracking/libs/CMakeLists.txt:44 changes:
// sm_30 -> sm_100a).  It is driven the way the reference drives it in
// src/tests/unit-tests/signal-processing-blocks/tracking/gpu_multicorrelator_test.cc:52-140: zero-copy mapped pinned host
// buffers (cudaHostAllocMapped | cudaHostAllocWriteCombined), one cuda_multicorrelator object and one std::thread per channel,
// every Carrier_wipeoff_multicorrelator_resampler_cuda call = one launch (128 blocks x 64 threads) + cudaStreamSynchronize.
// Parameters of the reference test: carrier step 0.1 rad, code step 0.3 chips, rem code 0.4 chips, taps +-0.5 chip (:110-114),
// plus the product's own shape (N = 25000, 11 taps) on request.
//
// usage: legacy_cuda_bench <n_samples> <n_taps> <n_threads> <iters>   -> one JSON line on stdout
// The process never calls free_cuda() before it has printed (free_cuda ends in cudaDeviceReset, cuda_multicorrelator.cu:380).
#include "cuda_multicorrelator.h"
#include <chrono>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

static void run(cuda_multicorrelator* c, float rem_carr, float step, float code_step, float rem_code, int n, int taps, int iters)
{
    for (int k = 0; k < iters; k++) c->Carrier_wipeoff_multicorrelator_resampler_cuda(rem_carr, step, code_step, rem_code, n, taps);
}

int main(int argc, char** argv)
{
    const int n = argc > 1 ? atoi(argv[1]) : 8192;
    const int taps = argc > 2 ? atoi(argv[2]) : 3;
    const int n_threads = argc > 3 ? atoi(argv[3]) : 1;
    const int iters = argc > 4 ? atoi(argv[4]) : 1000;
    const int L = 1023;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        {
            printf("{\"legacy_cuda\": \"unavailable\", \"why\": \"no CUDA device\"}\n");
            return 0;
        }
    cudaSetDevice(0);
    cudaSetDeviceFlags(cudaDeviceMapHost);
    std::complex<float>*code, *in, *outs;
    float* shifts;
    cudaHostAlloc(reinterpret_cast<void**>(&code), L * sizeof(std::complex<float>), cudaHostAllocMapped | cudaHostAllocWriteCombined);
    cudaHostAlloc(reinterpret_cast<void**>(&shifts), taps * sizeof(float), cudaHostAllocMapped | cudaHostAllocWriteCombined);
    cudaHostAlloc(reinterpret_cast<void**>(&in), 2 * (size_t)n * sizeof(std::complex<float>), cudaHostAllocMapped | cudaHostAllocWriteCombined);
    cudaHostAlloc(reinterpret_cast<void**>(&outs), (size_t)n_threads * taps * sizeof(std::complex<float>), cudaHostAllocMapped);
    srand(1);
    std::vector<std::complex<float>> code_h(L), in_h(2 * (size_t)n);
    for (int i = 0; i < L; i++) code_h[i] = code[i] = std::complex<float>((rand() & 1) ? 1.0f : -1.0f, 0.0f);
    for (int i = 0; i < 2 * n; i++)
        in_h[i] = in[i] = std::complex<float>(static_cast<float>(rand()) / static_cast<float>(RAND_MAX), static_cast<float>(rand()) / static_cast<float>(RAND_MAX));
    for (int t = 0; t < taps; t++) shifts[t] = taps == 3 ? (t - 1) * 0.5f : (t - (taps - 1) / 2.0f) * 0.1f;
    const float rem_carr = 0.0f, step = 0.1f, code_step = n >= 20000 ? 1.023e6f / 25e6f : 0.3f, rem_code = 0.4f;

    std::vector<cuda_multicorrelator*> pool(n_threads);
    for (int k = 0; k < n_threads; k++)
        {
            pool[k] = new cuda_multicorrelator();
            pool[k]->init_cuda_integrated_resampler(2 * n, L, taps);
            pool[k]->set_local_code_and_taps(L, code, shifts, taps);
            pool[k]->set_input_output_vectors(outs + (size_t)k * taps, in);
        }
    cudaDeviceSynchronize();
    cudaError_t e0 = cudaGetLastError();
    // warm-up, then the timed region (host wall clock around launch+sync per call, as the reference test measures)
    run(pool[0], rem_carr, step, code_step, rem_code, n, taps, 20);
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int k = 0; k < n_threads; k++) th.emplace_back(run, pool[k], rem_carr, step, code_step, rem_code, n, taps, iters);
    for (auto& t : th) t.join();
    const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    cudaError_t e1 = cudaGetLastError();
    // accuracy against the float64 evaluation of the correlator law (floor(step*n + shift - rem) mod L, exp(-j(phi + n*step)))
    std::vector<std::complex<double>> want(taps);
    double scale = 0.0, err = 0.0;
    for (int t = 0; t < taps; t++)
        {
            std::complex<double> acc(0, 0);
            for (int i = 0; i < n; i++)
                {
                    long idx = (long)std::floor((double)code_step * i + (double)shifts[t] - (double)rem_code) % L;
                    if (idx < 0) idx += L;
                    const double ph = -((double)rem_carr + (double)i * (double)step);
                    acc += std::complex<double>(in_h[i]) * std::complex<double>(std::cos(ph), std::sin(ph)) * (double)code_h[idx].real();
                }
            want[t] = acc;
            scale = std::max(scale, std::abs(acc));
        }
    for (int t = 0; t < taps; t++) err = std::max(err, std::abs(std::complex<double>(outs[t]) - want[t]));
    const double calls = (double)n_threads * iters;
    printf("{\"legacy_cuda\": \"cuda_multicorrelator.cu unmodified, sm_100a\", \"n_samples\": %d, \"n_taps\": %d, \"threads\": %d, \"iters\": %d, "
           "\"seconds\": %.6f, \"us_per_call\": %.3f, \"channel_msps\": %.3f, \"rel_err_vs_f64\": %.3e, \"cuda_error_before\": \"%s\", \"cuda_error_after\": \"%s\", "
           "\"launch\": \"128 blocks x 64 threads per call + cudaStreamSynchronize, zero-copy mapped host buffers\"}\n",
        n, taps, n_threads, iters, secs, secs / calls * 1e6 * n_threads, calls * n / secs / 1e6, err / scale, cudaGetErrorString(e0), cudaGetErrorString(e1));
    fflush(stdout);
    return 0;
}
