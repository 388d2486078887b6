// gp_tc.cu — config C3 on the tensor cores: the fp32 tier (north_star: 1e-5) of the GP classifier of gp.cu.
//
// Same model, plan, prepare (fp64 Cholesky), begin / finish stages as gp.cu (included below with GP_TC defined); the
// two triangular products of a gradient evaluation,  G = L W  and  D = L^T R,  run on the 5th-generation tensor cores:
//   * every operand is split into two fp16 terms, x = hi + lo / 2^11 (22 significand bits, the residual scaled so
//     that it keeps 11 bits of its own), and a product is three tcgen05.mma terms: hi*hi into one fp32 accumulator
//     in TMEM, hi*lo + lo*hi into a second one that the epilogue adds scaled by 2^-11 (lo*lo, 2^-22, is dropped);
//   * L and L^T are split once per data set (gp_tc_split_L); W and R are written as split-fp16 by the epilogue of
//     the product (or the begin / finish stage) that produces them, so no separate conversion pass runs per step;
//   * the split operands live in HBM as the shared-memory images of their 32 k x 128 tiles (canonical MN-major
//     SWIZZLE_128B layout of tcgen05, see tc_tile_off in gp.cu), so staging a k step is two linear 16 KB bulk
//     copies (cp.async.bulk + mbarrier complete_tx) issued by ONE thread: TMA-class staging without a tensor map;
//   * one CTA = one 128 x 128 output tile; warp 5 lane 0 issues the bulk copies, warp 4 lane 0 the MMAs, three 32 KB
//     stages (two CTAs per SM: one's epilogue overlaps the other's main loop) hand over through mbarriers (full: expect_tx bytes, empty: tcgen05.commit);
//   * the same four warps then read the accumulator with tcgen05.ld (thread = output row) and run the fused epilogue
//     of gp.cu (r = z - sigmoid(g) / the leapfrog update), writing fp64 state and the split-fp16 operand of the next
//     product.  Row tiles are issued longest-first (the k range of a tile is triangular).
#define GP_TC 1
#include "gp.cu"

#define TC_BK 32
#define TC_STAGES 3
#define TC_STAGE_BYTES 32768          // 2 x (32 k x 128 m) fp16 A + 2 x (32 k x 128 chains) fp16 B
#define TC_SPLIT_BYTES 8192
#define TC_TMEM_COLS 256              // two 128 x 128 fp32 accumulators
#define TC_EPI_WARPS 8                // epilogue warps (4 or 8); + 1 MMA warp + 1 producer warp
#define TC_EPI_ROWS (128 / TC_EPI_WARPS)
#define TC_THREADS (32 * (TC_EPI_WARPS + 2))

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
// instruction descriptor (InstrDescriptor): D = f32, A = B = f16 (format 0), A and B MN-major, N = 128, M = 128
#define TC_IDESC ((1u << 4) | (0u << 7) | (0u << 10) | (1u << 15) | (1u << 16) | ((128u >> 3) << 17) | ((128u >> 4) << 24))

// ---- operand staging: linear bulk copies of tile images --------------------------------------------------------------
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("{\n.reg .b64 st;\nmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void* src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(bar)
               : "memory");
}

__device__ __forceinline__ float sigmoidf_(float g) { return 1.f / (1.f + __expf(-g)); }
__device__ __forceinline__ float log1p_expf_(float v) { return v > 0.f ? v + log1pf(__expf(-v)) : log1pf(__expf(v)); }

template <int MODE>
__device__ __forceinline__ void gp_tc_gemm(const StkArgs& a) {
  extern __shared__ unsigned char smraw[];
  unsigned char* smem = (unsigned char*)(((unsigned long long)smraw + 1023) & ~1023ULL);
  unsigned long long* bars = (unsigned long long*)(smem + TC_STAGES * TC_STAGE_BYTES);   // full[4], empty[4], accum
  unsigned* tmem_slot = (unsigned*)(bars + 16);
  const long long n = a.flen[F_X], ld = ld_of(n), C = a.n_chains;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int nt = gridDim.y;
  const int I = MODE == 1 ? nt - 1 - (int)blockIdx.y : (int)blockIdx.y;
  const long long m0 = (long long)I * 128, c0 = (long long)blockIdx.x * 128;
  // A operand stored [k][m]: product 1 (A = L) reads the split of L^T's rows, product 2 (A = L^T) the split of L's
  const tc_t* Aop = (const tc_t*)a.field[MODE == 1 ? F_LTA : F_LA];
  const tc_t* Bop = (const tc_t*)a.field[MODE == 1 ? F_WB : F_RB];
  const long long kbeg = MODE == 1 ? 0 : m0, kend = MODE == 1 ? m0 + 128 : ld;
#ifdef TC_DIAG_K1         /* diagnostic build: one k step per tile = the fixed cost of a tile (set-up + epilogue) */
  const int ktiles = 1;
#else
  const int ktiles = (int)((kend - kbeg) / TC_BK);
#endif
  const long long ncbA = ld >> 7, ncbB = (C + 127) >> 7;
  int* info = (int*)a.field[F_INFO];
  const unsigned full0 = smem_u32(bars), empty0 = smem_u32(bars + TC_STAGES), accum = smem_u32(bars + 2 * TC_STAGES);

  if (tid == 0) {
    for (int i = 0; i < TC_STAGES; ++i) { mbar_init(full0 + 8 * i, 1); mbar_init(empty0 + 8 * i, 1); }
    mbar_init(accum, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == TC_EPI_WARPS) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(tmem_slot)), "n"(TC_TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const unsigned tmem = *tmem_slot;

  if (tid == 32 * (TC_EPI_WARPS + 1)) {
    // ===== producer (one thread): two 16 KB bulk copies per stage =====
    const tc_t* ga = Aop + ((kbeg >> 5) * ncbA + I) * (TC_NS * 4096);
    const tc_t* gb = Bop + ((kbeg >> 5) * ncbB + blockIdx.x) * (TC_NS * 4096);
    for (int t = 0; t < ktiles; ++t) {
      const int st = t % TC_STAGES;
      if (t >= TC_STAGES && !mbar_wait(empty0 + 8 * st, ((t / TC_STAGES) & 1) ^ 1)) info[1] = 1;
      const unsigned dst = smem_u32(smem + st * TC_STAGE_BYTES), bar = full0 + 8 * st;
      mbar_expect_tx(bar, TC_STAGE_BYTES);
      bulk_g2s(dst, ga, TC_NS * TC_SPLIT_BYTES, bar);
      bulk_g2s(dst + TC_NS * TC_SPLIT_BYTES, gb, TC_NS * TC_SPLIT_BYTES, bar);
      ga += ncbA * (TC_NS * 4096); gb += ncbB * (TC_NS * 4096);
    }
  } else if (tid == 32 * TC_EPI_WARPS) {
    // ===== MMA issuer (one thread) =====
    for (int t = 0; t < ktiles; ++t) {
      const int st = t % TC_STAGES;
      if (!mbar_wait(full0 + 8 * st, (t / TC_STAGES) & 1)) info[1] = 2;
      tc_fence_after();
      const unsigned sa = smem_u32(smem + st * TC_STAGE_BYTES), sb = sa + TC_NS * TC_SPLIT_BYTES;
      const unsigned long long a_hi = smem_desc(sa, 4096, 1024), a_lo = smem_desc(sa + TC_SPLIT_BYTES, 4096, 1024);
      const unsigned long long b_hi = smem_desc(sb, 4096, 1024), b_lo = smem_desc(sb + TC_SPLIT_BYTES, 4096, 1024);
#pragma unroll
      for (int k2 = 0; k2 < TC_BK / 16; ++k2) {   // UMMA_K = 16 = two 8-k groups (2048 B) of either operand
        const unsigned long long adv = (unsigned long long)(k2 * 128);
        const unsigned acc = (t | k2) != 0;
        tc_mma(tmem, a_hi + adv, b_hi + adv, TC_IDESC, acc);              // accumulator 0: hi * hi
#ifndef TC_DIAG_ONE_MMA   /* diagnostic build: only the hi * hi term (is the main loop bound by operand delivery or by MMA issue?) */
        tc_mma(tmem + 128, a_hi + adv, b_lo + adv, TC_IDESC, acc);        // accumulator 1: hi * lo + lo * hi
        tc_mma(tmem + 128, a_lo + adv, b_hi + adv, TC_IDESC, 1u);
#endif
      }
      tc_commit(empty0 + 8 * st);
    }
    tc_commit(accum);
  }

  // ===== epilogue (warps 0 .. TC_EPI_WARPS-1; the tile epilogue is a latency-bound instruction stream: measured with a one-k-step
  // build, 4 warps spent 49-58 us of a 91 us launch here, so it runs on 8 warps of 16 rows) =====
  // phase A: thread = accumulator row (TMEM lane): tcgen05.ld -> shared tile T[128][132] (the operand stages are free);
  // phase B: warp = 32 rows, lane = 4 consecutive chains: every global access of a warp is one contiguous run
  if (warp < TC_EPI_WARPS) {
    if (!mbar_wait(accum, 0)) info[1] = 3;
    tc_fence_after();
    float* T = (float*)smem;
    double* sc = (double*)(smem + 69632);         // [2][TC_EPI_WARPS][128] column sums of the epilogue warps
    {
      // a warp reaches the TMEM lanes 32 (warp % 4) ..: warps w and w + 4 share a lane quarter and split its columns
      const unsigned trow = tmem + ((unsigned)((warp & 3) * 32) << 16);
      const int trw = (warp & 3) * 32 + (tid & 31);              // accumulator row of this thread
      for (int j = (warp >> 2) * (16 / TC_EPI_WARPS); j < ((warp >> 2) + 1) * (16 / TC_EPI_WARPS); ++j) {
        float v[32], x[32];
        tc_ld32(trow + 32 * j, v);
        tc_ld32(trow + 128 + 32 * j, x);
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = fmaf(x[i], (float)(1.0 / TC_LO_SCALE), v[i]);
#pragma unroll
        for (int q = 0; q < 8; ++q)
          *reinterpret_cast<float4*>(T + trw * 132 + 32 * j + 4 * q) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
      }
    }
    tc_fence_before();
    asm volatile("bar.sync 1, %0;\n" ::"n"(32 * TC_EPI_WARPS) : "memory");
    const int lane = tid & 31;
    const long long cc = c0 + 4 * lane;
    const bool last = a.iter == a.n_iter - 1;
    const bool sums = MODE == 1 ? (a.stage_kind != 2 || last) : (MODE == 3 && last);
    const bool full4 = (C & 3) == 0 && cc + 4 <= C;
    const double eps = a.par[P_EPS];
    double* red = (double*)a.field[F_RED];
    double cs0[4] = {0.0, 0.0, 0.0, 0.0}, cs1[4] = {0.0, 0.0, 0.0, 0.0};
    const long long mw = m0 + warp * TC_EPI_ROWS;                // first row of this warp
    const int nrows = n - mw >= TC_EPI_ROWS ? TC_EPI_ROWS : (n > mw ? (int)(n - mw) : 0);
    const int zlane = (MODE == 1 && lane < nrows) ? ((const int*)a.field[F_Z])[mw + lane] : 0;
    double* Wm = (double*)a.field[F_W];
    double* Pm = (double*)a.field[F_P];
    double* GR = (double*)a.field[F_GR];
    tc_t* OB = (tc_t*)a.field[MODE == 1 ? F_RB : F_WB];       // split operand this epilogue produces
    const bool wr_gr = MODE == 2 || (MODE == 3 && last), wr_w = MODE == 3 && !last;
    constexpr int RBATCH = MODE == 3 ? 2 : 4;                    // the leapfrog epilogue holds w and p rows in registers: 102 per thread
    for (int rb = 0; rb < nrows; rb += RBATCH) {                 // RBATCH rows per batch: all global loads first
      double wv[RBATCH][4], pv[RBATCH][4];
      if (MODE != 1) {
#pragma unroll
        for (int u = 0; u < RBATCH; ++u) {
          const long long o = (mw + rb + u) * C + cc;
          if (rb + u < nrows) {
            if (full4) {
              const double2 w0 = *reinterpret_cast<const double2*>(Wm + o), w1 = *reinterpret_cast<const double2*>(Wm + o + 2);
              wv[u][0] = w0.x; wv[u][1] = w0.y; wv[u][2] = w1.x; wv[u][3] = w1.y;
              if (MODE == 3) {
                const double2 p0 = *reinterpret_cast<const double2*>(Pm + o), p1 = *reinterpret_cast<const double2*>(Pm + o + 2);
                pv[u][0] = p0.x; pv[u][1] = p0.y; pv[u][2] = p1.x; pv[u][3] = p1.y;
              }
            } else {
              for (int e = 0; e < 4; ++e) { wv[u][e] = cc + e < C ? Wm[o + e] : 0.0; pv[u][e] = (MODE == 3 && cc + e < C) ? Pm[o + e] : 0.0; }
            }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < RBATCH; ++u) {
        const int zi = __shfl_sync(0xffffffffu, zlane, (rb + u) & 31);
        if (rb + u >= nrows) continue;
        const long long m = mw + rb + u;
        const float4 g4 = *reinterpret_cast<const float4*>(T + (warp * TC_EPI_ROWS + rb + u) * 132 + 4 * lane);
        const float g[4] = {g4.x, g4.y, g4.z, g4.w};
        const long long o = m * C + cc;
        __align__(8) tc_t h[4], lo[4];
        double d[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (MODE == 1) {
            split2f((float)zi - sigmoidf_(g[e]), h[e], lo[e]);
            if (sums && cc + e < C) cs0[e] += (double)(zi ? -log1p_expf_(-g[e]) : -log1p_expf_(g[e]));
          } else {
            d[e] = (double)g[e] - wv[u][e];
            if (MODE == 3) {
              pv[u][e] += 0.5 * eps * d[e];
              if (!last) { pv[u][e] += 0.5 * eps * d[e]; wv[u][e] += eps * pv[u][e]; split2(wv[u][e], h[e], lo[e]); }
              else if (cc + e < C) { cs0[e] += pv[u][e] * pv[u][e]; cs1[e] += wv[u][e] * wv[u][e]; }
            }
          }
        }
        if ((MODE == 1 || wr_w) && cc < C) {   // 4 chains = half a 16-byte chunk of the tile image; chains >= C stay zero
          const long long ot = tc_tile_off(m, cc, ncbB);
          if (!full4) for (int e = 0; e < 4; ++e) if (cc + e >= C) { h[e] = lo[e] = __float2half(0.f); }
          *reinterpret_cast<uint2*>(OB + ot) = *reinterpret_cast<const uint2*>(h);
          *reinterpret_cast<uint2*>(OB + ot + TC_SPLIT_ELEMS) = *reinterpret_cast<const uint2*>(lo);
        }
        if (MODE != 1) {
          if (full4) {
            if (wr_gr) { *reinterpret_cast<double2*>(GR + o) = make_double2(d[0], d[1]); *reinterpret_cast<double2*>(GR + o + 2) = make_double2(d[2], d[3]); }
            if (MODE == 2 && a.out_grad) { *reinterpret_cast<double2*>(a.out_grad + o) = make_double2(d[0], d[1]); *reinterpret_cast<double2*>(a.out_grad + o + 2) = make_double2(d[2], d[3]); }
            if (MODE == 3) { *reinterpret_cast<double2*>(Pm + o) = make_double2(pv[u][0], pv[u][1]); *reinterpret_cast<double2*>(Pm + o + 2) = make_double2(pv[u][2], pv[u][3]); }
            if (wr_w) { *reinterpret_cast<double2*>(Wm + o) = make_double2(wv[u][0], wv[u][1]); *reinterpret_cast<double2*>(Wm + o + 2) = make_double2(wv[u][2], wv[u][3]); }
          } else {
            for (int e = 0; e < 4; ++e)
              if (cc + e < C) {
                if (wr_gr) GR[o + e] = d[e];
                if (MODE == 2 && a.out_grad) a.out_grad[o + e] = d[e];
                if (MODE == 3) Pm[o + e] = pv[u][e];
                if (wr_w) Wm[o + e] = wv[u][e];
              }
          }
        }
      }
    }
    if (sums) {   // column sums: rows in order inside a warp, then the four warps in order
#pragma unroll
      for (int e = 0; e < 4; ++e) { sc[warp * 128 + 4 * lane + e] = cs0[e]; sc[TC_EPI_WARPS * 128 + warp * 128 + 4 * lane + e] = cs1[e]; }
      asm volatile("bar.sync 1, %0;\n" ::"n"(32 * TC_EPI_WARPS) : "memory");
      if (tid < 128 && c0 + tid < C) {
        double t0 = 0.0, t1 = 0.0;
#pragma unroll
        for (int w8 = 0; w8 < TC_EPI_WARPS; ++w8) { t0 += sc[w8 * 128 + tid]; t1 += sc[TC_EPI_WARPS * 128 + w8 * 128 + tid]; }
        if (MODE == 1) red[((long long)RED_LIK * RED_TILES + I) * C + c0 + tid] = t0;
        else { red[((long long)RED_KE * RED_TILES + I) * C + c0 + tid] = t0; red[((long long)RED_WW * RED_TILES + I) * C + c0 + tid] = t1; }
      }
    }
  }
  __syncthreads();
  if (warp == TC_EPI_WARPS) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "n"(TC_TMEM_COLS));
}

extern "C" __global__ void __launch_bounds__(TC_THREADS, 2) gp_tc_gemm1(const __grid_constant__ StkArgs a) { gp_tc_gemm<1>(a); }
extern "C" __global__ void __launch_bounds__(TC_THREADS, 2) gp_tc_gemm2_grad(const __grid_constant__ StkArgs a) { gp_tc_gemm<2>(a); }
extern "C" __global__ void __launch_bounds__(TC_THREADS, 2) gp_tc_gemm2_leap(const __grid_constant__ StkArgs a) { gp_tc_gemm<3>(a); }

// ---- split stages -------------------------------------------------------------------------------------------------
// once per data set, both A operands as tile images of [k][m] matrices:
//   Lta (product 1, A = L):   element (k, m) = L[m][k] = Lt[k][m];    La (product 2, A = L^T): element (k, m) = L[k][m]
// the identity padding of Lk / Lt (rows >= n) is dropped: padded rows and columns of the operands are zero
extern "C" __global__ void __launch_bounds__(256) gp_tc_split_L(const __grid_constant__ StkArgs a) {
  const long long n = a.flen[F_X], ld = ld_of(n), tot = ld * ld;
  const double* Lk = (const double*)a.field[F_LK];
  const double* Lt = (const double*)a.field[F_LT];
  tc_t* La = (tc_t*)a.field[F_LA];
  tc_t* Lta = (tc_t*)a.field[F_LTA];
  const long long stride = (long long)gridDim.x * gridDim.y * 256;
  for (long long i = ((long long)blockIdx.y * gridDim.x + blockIdx.x) * 256 + threadIdx.x; i < tot; i += stride) {
    const long long k = i / ld, m = i % ld;
    const bool in = k < n && m < n;
    split_store_t(La, k, m, ld >> 7, in ? Lk[i] : 0.0);
    split_store_t(Lta, k, m, ld >> 7, in ? Lt[i] : 0.0);
  }
}

// wb = split(w) (after an upload of w / simulate; inside a trajectory the producing epilogue writes wb itself)
extern "C" __global__ void __launch_bounds__(128) gp_tc_split_w(const __grid_constant__ StkArgs a) {
  const long long c = (long long)blockIdx.x * 128 + threadIdx.x;
  if (c >= a.n_chains) return;
  const long long n = a.flen[F_X], C = a.n_chains;
  const double* W = (const double*)a.field[F_W];
  const long long j0 = (long long)blockIdx.y * 128, j1 = j0 + 128 < n ? j0 + 128 : n;
  for (long long j = j0; j < j1; ++j) split_store_t((tc_t*)a.field[F_WB], j, c, (C + 127) >> 7, W[j * C + c]);
}
