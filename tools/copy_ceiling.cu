// The host-side ceiling of the end-to-end path: what this box moves with NO kernels running.
//   nvcc -O2 -o tools/_bisect/copy_ceiling tools/copy_ceiling.cu && tools/_bisect/copy_ceiling [max_devices] > profiles/copy_ceiling_r2.json
// For N = 1, 2, 4, 8 GPUs at once, per GPU: 4 GiB pinned host -> device and 2.1 GiB device -> pinned host (the byte volumes
// of one bench step per GPU), queued asynchronously in 64 MiB pieces on two streams, both directions at once, each
// direction alone as well; best of 3.  One process, like the product's multi-device call.
#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)
int main(int argc, char **argv) {
    int ndev = 0; CK(cudaGetDeviceCount(&ndev));
    int maxdev = argc > 1 ? atoi(argv[1]) : 8; if (maxdev > ndev) maxdev = ndev;
    const size_t PIECE = 64ull << 20, H2D = 4ull << 30, D2H = (size_t)(2.1 * (1ull << 30)) / PIECE * PIECE;
    struct dev { char *h_in, *h_out, *d_in, *d_out; cudaStream_t si, so; };
    std::vector<dev> D(maxdev);
    for (int d = 0; d < maxdev; ++d) {
        CK(cudaSetDevice(d));
        CK(cudaHostAlloc(&D[d].h_in, 4 * PIECE, cudaHostAllocPortable)); CK(cudaHostAlloc(&D[d].h_out, 4 * PIECE, cudaHostAllocPortable));
        CK(cudaMalloc(&D[d].d_in, 4 * PIECE)); CK(cudaMalloc(&D[d].d_out, 4 * PIECE));
        CK(cudaStreamCreateWithFlags(&D[d].si, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&D[d].so, cudaStreamNonBlocking));
    }
    auto run = [&](int n, bool in, bool out) {
        double best = 1e9;
        for (int rep = 0; rep < 3; ++rep) {
            const double t0 = now();
            for (size_t o = 0; o < H2D; o += PIECE)          /* interleave the devices while queueing */
                for (int d = 0; d < n; ++d) {
                    CK(cudaSetDevice(d));
                    const size_t k = (o / PIECE) & 3;
                    if (in) CK(cudaMemcpyAsync(D[d].d_in + k * PIECE, D[d].h_in + k * PIECE, PIECE, cudaMemcpyHostToDevice, D[d].si));
                    if (out && o < D2H) CK(cudaMemcpyAsync(D[d].h_out + k * PIECE, D[d].d_out + k * PIECE, PIECE, cudaMemcpyDeviceToHost, D[d].so));
                }
            for (int d = 0; d < n; ++d) { CK(cudaSetDevice(d)); CK(cudaStreamSynchronize(D[d].si)); CK(cudaStreamSynchronize(D[d].so)); }
            const double t = now() - t0;
            if (t < best) best = t;
        }
        return best;
    };
    printf("{\"h2d_bytes_per_gpu\": %zu, \"d2h_bytes_per_gpu\": %zu, \"piece\": %zu, \"runs\": [", H2D, D2H, PIECE);
    bool first = true;
    for (int n = 1; n <= maxdev; n *= 2) {
        const double tb = run(n, true, true), ti = run(n, true, false), to = run(n, false, true);
        printf("%s{\"n_gpus\": %d, \"both_s\": %.5f, \"both_input_GBps\": %.2f, \"both_dma_GBps\": %.2f, \"h2d_only_GBps\": %.2f, \"d2h_only_GBps\": %.2f}",
               first ? "" : ", ", n, tb, n * H2D / tb / 1e9, n * (double)(H2D + D2H) / tb / 1e9, n * H2D / ti / 1e9, n * D2H / to / 1e9);
        first = false;
    }
    printf("]}\n");
    return 0;
}
