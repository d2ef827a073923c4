// Diagnostic (not part of the product): how fast can 32-byte records be read at random positions?
// Variants of the load instruction / path; prints ms and effective GB/s (useful bytes: 32 B read + 32 B written per record).
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <random>
#include <cuda_runtime.h>

struct __align__(32) Rec { uint64_t a, b, c, d; };

__device__ __forceinline__ Rec ld_nc256(const Rec *p) {
  Rec r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r.a), "=l"(r.b), "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}
__device__ __forceinline__ Rec ld_2x128(const Rec *p) {
  Rec r;
  asm volatile("ld.global.nc.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
  asm volatile("ld.global.nc.v2.u64 {%0,%1}, [%2+16];" : "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}
__device__ __forceinline__ Rec ld_cg(const Rec *p) {
  Rec r;
  asm volatile("ld.global.cg.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
  asm volatile("ld.global.cg.v2.u64 {%0,%1}, [%2+16];" : "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}
__device__ __forceinline__ Rec ld_cv(const Rec *p) {   // volatile: bypass caches where possible
  Rec r;
  asm volatile("ld.global.cv.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
  asm volatile("ld.global.cv.v2.u64 {%0,%1}, [%2+16];" : "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}
__device__ __forceinline__ Rec ld_lu(const Rec *p) {   // last use
  Rec r;
  asm volatile("ld.global.lu.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
  asm volatile("ld.global.lu.v2.u64 {%0,%1}, [%2+16];" : "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}
__device__ __forceinline__ Rec ld_evict_first(const Rec *p) {
  Rec r;
  asm volatile("ld.global.L1::no_allocate.L2::evict_first.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r.a), "=l"(r.b), "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}

__device__ __forceinline__ Rec ld_l2_64(const Rec *p) {
  Rec r;
  asm volatile("ld.global.L2::64B.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
  asm volatile("ld.global.L2::64B.v2.u64 {%0,%1}, [%2+16];" : "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}
__device__ __forceinline__ Rec ld_l2_128(const Rec *p) {
  Rec r;
  asm volatile("ld.global.L2::128B.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
  asm volatile("ld.global.L2::128B.v2.u64 {%0,%1}, [%2+16];" : "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}
__device__ __forceinline__ Rec ld_l2_256(const Rec *p) {
  Rec r;
  asm volatile("ld.global.L2::256B.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
  asm volatile("ld.global.L2::256B.v2.u64 {%0,%1}, [%2+16];" : "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}

__device__ __forceinline__ Rec ld_nc_l2_64(const Rec *p) {
  Rec r;
  asm volatile("ld.global.nc.L2::64B.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
  asm volatile("ld.global.nc.L2::64B.v2.u64 {%0,%1}, [%2+16];" : "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}
__device__ __forceinline__ Rec ld_256_l2_64(const Rec *p) {
  Rec r;
  asm volatile("ld.global.L2::64B.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r.a), "=l"(r.b), "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}
__device__ __forceinline__ Rec ld_nc256_l2_64(const Rec *p) {
  Rec r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(r.a), "=l"(r.b), "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}
__device__ __forceinline__ Rec ld_plain2(const Rec *p) {
  Rec r;
  asm volatile("ld.global.v2.u64 {%0,%1}, [%2];" : "=l"(r.a), "=l"(r.b) : "l"(p));
  asm volatile("ld.global.v2.u64 {%0,%1}, [%2+16];" : "=l"(r.c), "=l"(r.d) : "l"(p));
  return r;
}

template <int V>
__device__ __forceinline__ Rec ld_variant(const Rec *p) {
  if (V == 10) return ld_nc_l2_64(p);
  if (V == 11) return ld_256_l2_64(p);
  if (V == 12) return ld_nc256_l2_64(p);
  if (V == 13) return ld_plain2(p);
  if (V == 0) return ld_nc256(p);
  if (V == 1) return ld_2x128(p);
  if (V == 2) return ld_cg(p);
  if (V == 3) return ld_cv(p);
  if (V == 4) return ld_lu(p);
  if (V == 5) return ld_evict_first(p);
  if (V == 6) return ld_l2_64(p);
  if (V == 7) return ld_l2_128(p);
  if (V == 8) return ld_l2_256(p);
  return *p;
}

template <int V, int ILP = 1>
__global__ void __launch_bounds__(256) k_gather(const Rec *__restrict__ in, const uint32_t *__restrict__ perm, size_t n, Rec *__restrict__ out) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + (ILP - 1) * stride < n; i += ILP * stride) {
    Rec r[ILP];
#pragma unroll
    for (int u = 0; u < ILP; ++u) r[u] = ld_variant<V>(in + perm[i + u * stride]);
#pragma unroll
    for (int u = 0; u < ILP; ++u) out[i + u * stride] = r[u];
  }
  for (; i < n; i += stride) out[i] = ld_variant<V>(in + perm[i]);
}

// TMA-style bulk copies: each thread pulls its 32-byte record into shared memory with cp.async.bulk, then writes it out
__global__ void __launch_bounds__(256) k_gather_bulk(const Rec *__restrict__ in, const uint32_t *__restrict__ perm, size_t n, Rec *__restrict__ out) {
  __shared__ __align__(128) Rec tile[256];
  __shared__ uint64_t bar;
  const unsigned tid = threadIdx.x;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)), "r"(1) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  uint32_t phase = 0;
  const size_t n_tiles = (n + 255) / 256;
  for (size_t t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const size_t i = t * 256 + tid;
    const uint32_t cnt = (uint32_t)min((size_t)256, n - t * 256);
    if (tid == 0)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)), "r"(cnt * 32u) : "memory");
    __syncthreads();
    if (i < n)
      asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], 32, [%2];" ::"r"(
                       (uint32_t)__cvta_generic_to_shared(&tile[tid])),
                   "l"(in + perm[i]), "r"((uint32_t)__cvta_generic_to_shared(&bar))
                   : "memory");
    uint32_t ok;
    do {
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"((uint32_t)__cvta_generic_to_shared(&bar)), "r"(phase) : "memory");
    } while (!ok);
    phase ^= 1;
    if (i < n) out[i] = tile[tid];
    __syncthreads();
  }
}

__global__ void k_copy(const Rec *__restrict__ in, size_t n, Rec *__restrict__ out) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = ld_nc256(in + i);
}

int main(int argc, char **argv) {
  const size_t n = argc > 1 ? strtoull(argv[1], 0, 10) : 10000000;
  int gran = argc > 2 ? atoi(argv[2]) : 0;
  if (gran) { cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, gran); printf("set fetch granularity %d: %s\n", gran, cudaGetErrorString(e)); }
  size_t g = 0; cudaDeviceGetLimit(&g, cudaLimitMaxL2FetchGranularity); printf("L2 fetch granularity limit: %zu\n", g);
  std::vector<uint32_t> perm(n);
  for (size_t i = 0; i < n; ++i) perm[i] = (uint32_t)i;
  std::mt19937_64 rng(1);
  std::shuffle(perm.begin(), perm.end(), rng);
  Rec *in, *out; uint32_t *dperm;
  cudaMalloc(&in, n * 32); cudaMalloc(&out, n * 32); cudaMalloc(&dperm, n * 4);
  cudaMemset(in, 1, n * 32);
  cudaMemcpy(dperm, perm.data(), n * 4, cudaMemcpyHostToDevice);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const int grid = 148 * 8;
  auto run = [&](const char *name, auto launch) {
    for (int w = 0; w < 2; ++w) launch();
    cudaEventRecord(a);
    for (int r = 0; r < 10; ++r) launch();
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); ms /= 10;
    printf("%-28s %.4f ms  %.0f GB/s useful  (%s)\n", name, ms, n * 64.0 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
  };
  run("sequential copy", [&] { k_copy<<<grid, 256>>>(in, n, out); });
  run("ld.nc.no_allocate.v4.u64", [&] { k_gather<0><<<grid, 256>>>(in, dperm, n, out); });
  run("2 x ld.nc.v2.u64", [&] { k_gather<1><<<grid, 256>>>(in, dperm, n, out); });
  run("2 x ld.cg.v2.u64", [&] { k_gather<2><<<grid, 256>>>(in, dperm, n, out); });
  run("2 x ld.cv.v2.u64", [&] { k_gather<3><<<grid, 256>>>(in, dperm, n, out); });
  run("2 x ld.lu.v2.u64", [&] { k_gather<4><<<grid, 256>>>(in, dperm, n, out); });
  run("ld.L2::evict_first.v4.u64", [&] { k_gather<5><<<grid, 256>>>(in, dperm, n, out); });
  run("plain struct load", [&] { k_gather<9><<<grid, 256>>>(in, dperm, n, out); });
  run("cp.async.bulk 32 B", [&] { k_gather_bulk<<<grid, 256>>>(in, dperm, n, out); });
  run("nc256, 4 in flight", [&] { k_gather<0, 4><<<grid, 256>>>(in, dperm, n, out); });
  run("nc256, 8 in flight", [&] { k_gather<0, 8><<<grid, 256>>>(in, dperm, n, out); });
  run("nc256, 16 in flight", [&] { k_gather<0, 16><<<grid, 256>>>(in, dperm, n, out); });
  run("L2::64B, 8 in flight", [&] { k_gather<6, 8><<<grid, 256>>>(in, dperm, n, out); });
  run("L2::128B, 8 in flight", [&] { k_gather<7, 8><<<grid, 256>>>(in, dperm, n, out); });
  run("L2::256B, 8 in flight", [&] { k_gather<8, 8><<<grid, 256>>>(in, dperm, n, out); });
  run("nc L2::64B 2x128, 8 in flight", [&] { k_gather<10, 8><<<grid, 256>>>(in, dperm, n, out); });
  run("L2::64B v4.u64, 8 in flight", [&] { k_gather<11, 8><<<grid, 256>>>(in, dperm, n, out); });
  run("nc.noalloc.L2::64B v4, 8", [&] { k_gather<12, 8><<<grid, 256>>>(in, dperm, n, out); });
  run("plain 2x128, 8 in flight", [&] { k_gather<13, 8><<<grid, 256>>>(in, dperm, n, out); });
  run("L2::64B 2x128, 1 in flight", [&] { k_gather<6, 1><<<grid, 256>>>(in, dperm, n, out); });
  run("evict_first, 8 in flight", [&] { k_gather<5, 8><<<grid, 256>>>(in, dperm, n, out); });
  run("nc256, 8 in flight, 4 CTA/SM", [&] { k_gather<0, 8><<<148 * 4, 256>>>(in, dperm, n, out); });
  return 0;
}
