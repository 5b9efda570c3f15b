// L2 probe ("an L2-bandwidth peak must be measured by the build"): the
// memory roofline the tracer actually lives under.  At 256^3 both grids are L2-resident, so the
// relevant peaks are L2 sector throughput for scattered 4-byte gathers and for float atomics
// (RED.ADD.F32), not HBM copy bandwidth.  Standalone binary, built by csrc/Makefile (`make probe`).
//
//   l2_probe [working-set MiB = 64] [iterations per thread = 256]
//
// Patterns (one warp-wide request each):
//   coalesced   32 lanes -> 32 consecutive floats (4 sectors)         the adjoint with interleave
//   cell8       8 lanes share an address, 4 distinct x-adjacent cells  8 rays per cell, no interleave
//   scattered   32 lanes -> 32 random sectors                         oblique / cone batches
//   hot32       every lane of every warp -> ONE address               a focal point (Luneburg lens)
//   hot1        lane 0 of every warp -> the same ONE address          ... after a warp-level pre-reduction
//   hot8x32     every warp -> one of 8 addresses, 32 lanes each       the 8 corners of a focal cell
// Output: one line per (op, pattern): G lane-ops/s, G sectors/s, GB/s at 32 B per sector.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

enum { COALESCED = 0, CELL8 = 1, SCATTERED = 2, HOT32 = 3, HOT1 = 4, HOT8X32 = 5 };

__device__ __forceinline__ uint32_t mix(uint32_t x) {  // xorshift-multiply hash
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

template <int PATTERN>
__device__ __forceinline__ uint32_t index_of(uint32_t warp, uint32_t lane, uint32_t it, uint32_t mask) {
  const uint32_t base = mix(warp * 0x9E3779B9u + it * 0x85EBCA6Bu);
  if (PATTERN == COALESCED) return ((base & ~31u) + lane) & mask;
  if (PATTERN == CELL8) return ((base & ~3u) + (lane >> 3)) & mask;
  if (PATTERN == HOT32 || PATTERN == HOT1) return 12345u & mask;
  if (PATTERN == HOT8X32) return (12345u + 1031u * (base & 7u)) & mask;
  return mix(base + lane * 0xC2B2AE35u) & mask;  // SCATTERED
}

template <int PATTERN>
__global__ void __launch_bounds__(256) gather_kernel(const float *__restrict__ buf, uint32_t mask, int iters, float *sink) {
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  float acc = 0.f;
#pragma unroll 8
  for (int it = 0; it < iters; ++it) acc += __ldg(buf + index_of<PATTERN>(warp, lane, it, mask));
  if (acc == 123.456f) *sink = acc;
}

template <int PATTERN>
__global__ void __launch_bounds__(256) red_kernel(float *buf, uint32_t mask, int iters) {
  const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
#pragma unroll 8
  for (int it = 0; it < iters; ++it) {
    if (PATTERN == HOT1 && lane != 0) continue;
    atomicAdd(buf + index_of<PATTERN>(warp, lane, it, mask), 1.0f);
  }
}

static double sectors_per_request(int pattern) { return pattern == COALESCED ? 4.0 : pattern == SCATTERED ? 32.0 : 1.0; }

int main(int argc, char **argv) {
  const int mib = argc > 1 ? atoi(argv[1]) : 64;
  const int iters = argc > 2 ? atoi(argv[2]) : 256;
  size_t n = (size_t) mib << 18;  // floats
  size_t p2 = 1; while (p2 * 2 <= n) p2 *= 2;
  n = p2;
  const uint32_t mask = (uint32_t) (n - 1);
  float *buf, *sink;
  CK(cudaMalloc(&buf, n * sizeof(float)));
  CK(cudaMalloc(&sink, sizeof(float)));
  CK(cudaMemset(buf, 0, n * sizeof(float)));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  const int blocks = prop.multiProcessorCount * 8 * 4;  // 4 waves of 8 resident 256-thread CTAs per SM
  const double lanes = (double) blocks * 256 * iters;
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  printf("# l2_probe: %s, %d SMs, working set %zu MiB, %d CTAs x 256 threads x %d iterations\n", prop.name,
         prop.multiProcessorCount, n >> 18, blocks, iters);
  const char *pn[6] = {"coalesced", "cell8", "scattered", "hot32", "hot1", "hot8x32"};
  for (int op = 0; op < 2; ++op) {
    for (int pat = 0; pat < (op == 0 ? 3 : 6); ++pat) {
      float best = 1e30f;
      for (int rep = 0; rep < 4; ++rep) {  // rep 0 warms the L2
        CK(cudaEventRecord(a));
        if (op == 0) {
          if (pat == COALESCED) gather_kernel<COALESCED><<<blocks, 256>>>(buf, mask, iters, sink);
          else if (pat == CELL8) gather_kernel<CELL8><<<blocks, 256>>>(buf, mask, iters, sink);
          else gather_kernel<SCATTERED><<<blocks, 256>>>(buf, mask, iters, sink);
        } else {
          if (pat == COALESCED) red_kernel<COALESCED><<<blocks, 256>>>(buf, mask, iters);
          else if (pat == CELL8) red_kernel<CELL8><<<blocks, 256>>>(buf, mask, iters);
          else if (pat == SCATTERED) red_kernel<SCATTERED><<<blocks, 256>>>(buf, mask, iters);
          else if (pat == HOT32) red_kernel<HOT32><<<blocks, 256>>>(buf, mask, iters / 16);
          else if (pat == HOT1) red_kernel<HOT1><<<blocks, 256>>>(buf, mask, iters / 16);
          else red_kernel<HOT8X32><<<blocks, 256>>>(buf, mask, iters / 16);
        }
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        if (rep > 0 && ms < best) best = ms;
      }
      const double nl = pat >= HOT32 ? lanes / 16.0 / (pat == HOT1 ? 32.0 : 1.0) : lanes;  // hot patterns run iters / 16
      const double s = best * 1e-3, req = (pat >= HOT32 ? lanes / 16.0 : lanes) / 32.0, sect = req * sectors_per_request(pat);
      printf("%-6s %-10s %8.3f ms  %8.1f G lane-ops/s  %8.2f G requests/s  %8.1f G sectors/s  %8.1f GB/s (32 B sectors)\n",
             op == 0 ? "gather" : "red", pn[pat], best, nl / s * 1e-9, req / s * 1e-9, sect / s * 1e-9, sect * 32.0 / s * 1e-9);
    }
  }
  CK(cudaGetLastError());
  cudaFree(buf); cudaFree(sink);
  return 0;
}
