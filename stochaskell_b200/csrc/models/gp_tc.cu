// gp_tc.cu — config C3 on the tensor cores: the fp32 tier (north_star: 1e-5) of the GP classifier of gp.cu.
//
// Same model, plan, prepare (fp64 Cholesky), begin / finish stages as gp.cu (included below with GP_TC defined); the
// two triangular products of a gradient evaluation,  G = L W  and  D = L^T R,  run on the 5th-generation tensor cores:
//   * every operand is split into three bf16 terms, x = hi + mid + lo (24 significand bits), and a product is the
//     six tcgen05.mma terms hi*hi + hi*mid + mid*hi + mid*mid + hi*lo + lo*hi accumulated in fp32 in TMEM;
//   * L and L^T are split once per data set (gp_tc_split_L); W and R are written as split-bf16 by the epilogue of
//     the product (or the begin / finish stage) that produces them, so no separate conversion pass runs per step;
//   * one CTA = one 128 x 128 output tile; warps 0-3 stage the operand tiles with cp.async into the canonical
//     128-byte-swizzled layouts (A = L rows: K-major; B = W/R rows [k][chain]: MN-major), warp 4 lane 0 issues the
//     MMAs, two 96 KB stages hand over through mbarriers (full: 128 producer arrivals, empty: tcgen05.commit);
//   * the same four warps then read the accumulator with tcgen05.ld (thread = output row) and run the fused epilogue
//     of gp.cu (r = z - sigmoid(g) / the leapfrog update), writing fp64 state and the split-bf16 operand of the next
//     product.  Row tiles are issued longest-first (the k range of a tile is triangular).
#define GP_TC 1
#include "gp.cu"

#define TC_BK 64
#define TC_STAGE_BYTES 98304          // 3 x (128 x 64) bf16 A + 3 x (64 x 128) bf16 B
#define TC_SPLIT_BYTES 16384

// ---- PTX wrappers -----------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
  asm volatile("{\n.reg .b64 st;\nmbarrier.arrive.shared::cta.b64 st, [%0];\n}\n" ::"r"(bar) : "memory");
}
// bounded wait: a protocol error must not hang the GPU; on time-out the caller's `info[1]` is set and the kernel runs on
__device__ __forceinline__ bool mbar_wait(unsigned bar, unsigned parity) {
  for (unsigned it = 0; it < (1u << 24); ++it) {
    unsigned ok;
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return true;
  }
  return false;
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_commit(unsigned bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_mma(unsigned tmem_d, unsigned long long da, unsigned long long db, unsigned idesc, unsigned acc) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void tc_ld32(unsigned taddr, float* v) {
  unsigned* r = reinterpret_cast<unsigned*>(v);
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
}

// shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp SmemDescriptor): start >> 4 | LBO >> 4 << 16 |
// SBO >> 4 << 32 | version 1 << 46 | SWIZZLE_128B (2) << 61
__device__ __forceinline__ unsigned long long smem_desc(unsigned addr, unsigned lbo_bytes, unsigned sbo_bytes) {
  return (unsigned long long)((addr >> 4) & 0x3fff) | ((unsigned long long)((lbo_bytes >> 4) & 0x3fff) << 16) |
         ((unsigned long long)((sbo_bytes >> 4) & 0x3fff) << 32) | (1ULL << 46) | (2ULL << 61);
}
// instruction descriptor (InstrDescriptor): D = f32, A = B = bf16, A K-major, B MN-major, N = 128, M = 128
#define TC_IDESC ((1u << 4) | (1u << 7) | (1u << 10) | (0u << 15) | (1u << 16) | ((128u >> 3) << 17) | ((128u >> 4) << 24))

// ---- operand staging (warps 0-3) ---------------------------------------------------------------------------------
// A tile, K-major, 128 rows x 64 k, SWIZZLE_128B: byte (r, k) = (r/8) 1024 + (r%8) 128 + ((k/8) ^ (r%8)) 16 + (k%8) 2
// B tile, MN-major, 64 k x 128 columns, SWIZZLE_128B, two 64-column atoms 8192 B apart:
//   byte (k, c) = (c/64) 8192 + (k/8) 1024 + (k%8) 128 + (((c%64)/8) ^ (k%8)) 16 + (c%8) 2
__device__ __forceinline__ void tc_load_stage(unsigned char* st, const __nv_bfloat16* Aop, const __nv_bfloat16* Bop, long long ld,
                                              long long C, long long m0, long long c0, long long k0, int tid, bool vec) {
  const long long planeA = ld * ld, planeB = ld * C;
#pragma unroll
  for (int s = 0; s < 3; ++s) {
    unsigned char* as = st + s * TC_SPLIT_BYTES;
    unsigned char* bs = st + 3 * TC_SPLIT_BYTES + s * TC_SPLIT_BYTES;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int q = tid + 128 * u;
      {  // A: 128 rows x 8 chunks
        const int r = q >> 3, j = q & 7;
        cp_async16(as + (r >> 3) * 1024 + (r & 7) * 128 + ((j ^ (r & 7)) << 4), Aop + s * planeA + (m0 + r) * ld + k0 + j * 8, 16);
      }
      {  // B: 64 k x 16 chunks
        const int k = q >> 4, jj = q & 15, j = jj & 7;
        unsigned char* dst = bs + (jj >> 3) * 8192 + (k >> 3) * 1024 + (k & 7) * 128 + ((j ^ (k & 7)) << 4);
        const long long c = c0 + jj * 8;
        const __nv_bfloat16* src = Bop + s * planeB + (k0 + k) * C + c;
        if (vec) cp_async16(dst, c < C ? src : Bop, c < C ? 16 : 0);
        else {   // chain count not a multiple of 8: rows are not 16-byte aligned
          __nv_bfloat16* d = reinterpret_cast<__nv_bfloat16*>(dst);
          for (int e = 0; e < 8; ++e) d[e] = c + e < C ? src[e] : __float2bfloat16(0.f);
        }
      }
    }
  }
  cp_async_commit();
}

__device__ __forceinline__ float sigmoidf_(float g) { return 1.f / (1.f + __expf(-g)); }
__device__ __forceinline__ float log1p_expf_(float v) { return v > 0.f ? v + log1pf(__expf(-v)) : log1pf(__expf(v)); }

template <int MODE>
__device__ __forceinline__ void gp_tc_gemm(const StkArgs& a) {
  extern __shared__ unsigned char smraw[];
  unsigned char* smem = (unsigned char*)(((unsigned long long)smraw + 1023) & ~1023ULL);
  unsigned long long* bars = (unsigned long long*)(smem + 2 * TC_STAGE_BYTES);   // full[2], empty[2], accum
  unsigned* tmem_slot = (unsigned*)(bars + 8);
  const long long n = a.flen[F_X], ld = ld_of(n), C = a.n_chains;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int nt = gridDim.y;
  const int I = MODE == 1 ? nt - 1 - (int)blockIdx.y : (int)blockIdx.y;
  const long long m0 = (long long)I * 128, c0 = (long long)blockIdx.x * 128;
  const __nv_bfloat16* Aop = (const __nv_bfloat16*)a.field[MODE == 1 ? F_LA : F_LTA];
  const __nv_bfloat16* Bop = (const __nv_bfloat16*)a.field[MODE == 1 ? F_WB : F_RB];
  const long long kbeg = MODE == 1 ? 0 : m0, kend = MODE == 1 ? m0 + 128 : ld;
  const int ktiles = (int)((kend - kbeg) / TC_BK);
  const bool vec = (C & 7) == 0;
  int* info = (int*)a.field[F_INFO];
  const unsigned full0 = smem_u32(bars), empty0 = smem_u32(bars + 2), accum = smem_u32(bars + 4);

  if (tid == 0) {
    mbar_init(full0, 128); mbar_init(full0 + 8, 128);
    mbar_init(empty0, 1); mbar_init(empty0 + 8, 1);
    mbar_init(accum, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 4) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(tmem_slot)), "n"(128));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const unsigned tmem = *tmem_slot;

  if (warp < 4) {
    // ===== producers =====
    for (int t = 0; t < ktiles; ++t) {
      const int st = t & 1;
      if (t >= 2 && !mbar_wait(empty0 + 8 * st, ((t >> 1) & 1) ^ 1)) info[1] = 1;
      tc_load_stage(smem + st * TC_STAGE_BYTES, Aop, Bop, ld, C, m0, c0, kbeg + (long long)t * TC_BK, tid, vec);
      if (t >= 1) { cp_async_wait<1>(); fence_proxy_async(); mbar_arrive(full0 + 8 * ((t - 1) & 1)); }
    }
    cp_async_wait<0>(); fence_proxy_async(); mbar_arrive(full0 + 8 * ((ktiles - 1) & 1));
  } else if (tid == 128) {
    // ===== MMA issuer (one thread) =====
    for (int t = 0; t < ktiles; ++t) {
      const int st = t & 1;
      if (!mbar_wait(full0 + 8 * st, (t >> 1) & 1)) info[1] = 2;
      tc_fence_after();
      const unsigned sa = smem_u32(smem + st * TC_STAGE_BYTES), sb = sa + 3 * TC_SPLIT_BYTES;
      // (A split, B split): smallest terms first, hi*hi last
      const int pa[6] = {2, 0, 1, 1, 0, 0}, pb[6] = {0, 2, 1, 0, 1, 0};
#pragma unroll
      for (int q = 0; q < 6; ++q) {
        const unsigned long long da = smem_desc(sa + pa[q] * TC_SPLIT_BYTES, 16, 1024);
        const unsigned long long db = smem_desc(sb + pb[q] * TC_SPLIT_BYTES, 8192, 1024);
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4)   // UMMA_K = 16: A advances 32 B inside the swizzle row, B two 8-k groups
          tc_mma(tmem, da + (unsigned long long)(k4 * 2), db + (unsigned long long)(k4 * 128), TC_IDESC, (t | q | k4) != 0);
      }
      tc_commit(empty0 + 8 * st);
    }
    tc_commit(accum);
  }

  // ===== epilogue (warps 0-3; thread = output row, TMEM lane) =====
  if (warp < 4) {
    if (!mbar_wait(accum, 0)) info[1] = 3;
    tc_fence_after();
    const long long m = m0 + tid;
    const unsigned trow = tmem + ((unsigned)(warp * 32) << 16);
    double* sc0 = (double*)smem;                 // [128][33] column-sum scratch (operand stages are free now)
    double* sc1 = sc0 + 128 * 33;
    double* red = (double*)a.field[F_RED];
    const long long planeB = ld * C;
    const bool last = a.iter == a.n_iter - 1;
    const bool sums = MODE == 1 ? (a.stage_kind != 2 || last) : (MODE == 3 && last);
    const int zi = (MODE == 1 && m < n) ? ((const int*)a.field[F_Z])[m] : 0;
    const double eps = a.par[P_EPS];
    for (int j = 0; j < 4; ++j) {
      float v[32];
      tc_ld32(trow + 32 * j, v);
      const long long cb = c0 + 32 * j;
      double s0[32], s1[32];
      if (MODE == 1) {
        __nv_bfloat16* RB = (__nv_bfloat16*)a.field[F_RB];
        __align__(16) __nv_bfloat16 h[32], md[32], lo[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float g = v[i];
          const double rv = (double)((float)zi - sigmoidf_(g));
          split3(rv, h[i], md[i], lo[i]);
          if (sums) s0[i] = (m < n && cb + i < C) ? (double)(zi ? -log1p_expf_(-g) : -log1p_expf_(g)) : 0.0;
        }
        if (m < n) {
          const long long o = m * C + cb;
          if (vec && cb + 32 <= C) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              *reinterpret_cast<uint4*>(RB + o + 8 * q) = *reinterpret_cast<const uint4*>(h + 8 * q);
              *reinterpret_cast<uint4*>(RB + planeB + o + 8 * q) = *reinterpret_cast<const uint4*>(md + 8 * q);
              *reinterpret_cast<uint4*>(RB + 2 * planeB + o + 8 * q) = *reinterpret_cast<const uint4*>(lo + 8 * q);
            }
          } else {
            for (int i = 0; i < 32; ++i)
              if (cb + i < C) { RB[o + i] = h[i]; RB[planeB + o + i] = md[i]; RB[2 * planeB + o + i] = lo[i]; }
          }
        }
      } else {
        double* Wm = (double*)a.field[F_W];
        double* Pm = (double*)a.field[F_P];
        double* GR = (double*)a.field[F_GR];
        __nv_bfloat16* WB = (__nv_bfloat16*)a.field[F_WB];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          s0[i] = 0.0; s1[i] = 0.0;
          if (m < n && cb + i < C) {
            const long long o = m * C + cb + i;
            const double wv = Wm[o];
            const double d = (double)v[i] - wv;
            if (MODE == 2) { GR[o] = d; if (a.out_grad) a.out_grad[o] = d; }
            else {
              double pv = Pm[o];
              pv += 0.5 * eps * d;
              if (!last) {
                pv += 0.5 * eps * d;
                const double wn = wv + eps * pv;
                Wm[o] = wn;
                split3_store(WB, planeB, o, wn);
              } else { GR[o] = d; s0[i] = pv * pv; s1[i] = wv * wv; }
              Pm[o] = pv;
            }
          }
        }
      }
      if (sums) {   // column sums over the 128 rows of the tile, fixed order
        asm volatile("bar.sync 1, 128;\n" ::: "memory");
#pragma unroll
        for (int i = 0; i < 32; ++i) { sc0[tid * 33 + i] = s0[i]; if (MODE == 3) sc1[tid * 33 + i] = s1[i]; }
        asm volatile("bar.sync 1, 128;\n" ::: "memory");
        if (tid < 32 && cb + tid < C) {
          double t0 = 0.0, t1 = 0.0;
          for (int r = 0; r < 128; ++r) { t0 += sc0[r * 33 + tid]; if (MODE == 3) t1 += sc1[r * 33 + tid]; }
          if (MODE == 1) red[((long long)RED_LIK * RED_TILES + I) * C + cb + tid] = t0;
          else { red[((long long)RED_KE * RED_TILES + I) * C + cb + tid] = t0; red[((long long)RED_WW * RED_TILES + I) * C + cb + tid] = t1; }
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "n"(128));
}

extern "C" __global__ void __launch_bounds__(160, 1) gp_tc_gemm1(const __grid_constant__ StkArgs a) { gp_tc_gemm<1>(a); }
extern "C" __global__ void __launch_bounds__(160, 1) gp_tc_gemm2_grad(const __grid_constant__ StkArgs a) { gp_tc_gemm<2>(a); }
extern "C" __global__ void __launch_bounds__(160, 1) gp_tc_gemm2_leap(const __grid_constant__ StkArgs a) { gp_tc_gemm<3>(a); }

// ---- split stages -------------------------------------------------------------------------------------------------
// once per data set: La[s][m][k] = split(L[m][k]), Lta[s][m][k] = split(L^T[m][k])  (both K-major A operands)
extern "C" __global__ void __launch_bounds__(256) gp_tc_split_L(const __grid_constant__ StkArgs a) {
  const long long n = a.flen[F_X], ld = ld_of(n), tot = ld * ld;
  const double* Lk = (const double*)a.field[F_LK];
  const double* Lt = (const double*)a.field[F_LT];
  __nv_bfloat16* La = (__nv_bfloat16*)a.field[F_LA];
  __nv_bfloat16* Lta = (__nv_bfloat16*)a.field[F_LTA];
  const long long stride = (long long)gridDim.x * gridDim.y * 256;
  for (long long i = ((long long)blockIdx.y * gridDim.x + blockIdx.x) * 256 + threadIdx.x; i < tot; i += stride) {
    // the identity padding of Lk / Lt (rows >= n) is dropped: padded rows and columns of the operands are zero
    const long long r = i / ld, c = i % ld;
    const bool in = r < n && c < n;
    split3_store(La, tot, i, in ? Lk[i] : 0.0);
    split3_store(Lta, tot, i, in ? Lt[i] : 0.0);
  }
}

// wb = split(w) (after an upload of w / simulate; inside a trajectory the producing epilogue writes wb itself)
extern "C" __global__ void __launch_bounds__(128) gp_tc_split_w(const __grid_constant__ StkArgs a) {
  const long long c = (long long)blockIdx.x * 128 + threadIdx.x;
  if (c >= a.n_chains) return;
  const long long n = a.flen[F_X], C = a.n_chains;
  const double* W = (const double*)a.field[F_W];
  const long long j0 = (long long)blockIdx.y * 128, j1 = j0 + 128 < n ? j0 + 128 : n;
  for (long long j = j0; j < j1; ++j) split3_store((__nv_bfloat16*)a.field[F_WB], ld_of(n) * C, j * C + c, W[j * C + c]);
}
