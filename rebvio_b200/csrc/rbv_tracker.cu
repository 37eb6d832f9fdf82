// rbv_tracker.cu — edge tracker on sm_100a: distance field, keyline rotation, sigma_rho quantile, energy / Jacobian
// evaluation with on-device Levenberg-Marquardt, forward match, linearised 6-DoF solve, directed match, depth
// regularisation, per-keyline Kalman update.
//
// Replaces (reference paths relative to the reference root):
//   DistanceField::build                 rebvio/include/rebvio/core.hpp:37-59      -> k_field_build (packed atomicMin)
//   EdgeMap::rotateKeylines              rebvio/src/edge_map.cpp:58-71             -> k_rotate
//   EdgeMap::estimateQuantile            rebvio/src/edge_map.cpp:39-56             -> k_quantile
//   Core::tryVel/calculatefJ/testfk      rebvio/src/core.cpp:39-148                -> k_tryvel_fast (FAST) | k_tryvel_eval + k_tryvel_reduce (ORDERED)
//   Core::minimizeVel                    rebvio/src/core.cpp:150-189               -> LM step at the tail of k_tryvel_fast / k_tryvel_reduce
//   EdgeMap::forwardMatch                rebvio/src/edge_map.cpp:73-99             -> k_fwd_link + k_fwd_resolve
//   Core::extRotVel                      rebvio/src/core.cpp:191-261               -> k_extrot_fast (FAST) | k_extrot_rows + k_extrot_reduce (ORDERED)
//   EdgeMap::directedMatch/searchMatch   rebvio/src/edge_map.cpp:101-218           -> k_directed
//   EdgeMap::regularize1Iter             rebvio/src/edge_map.cpp:220-259           -> k_reg_eval + k_reg_apply
//   Core::updateInverseDepth(ARLU)       rebvio/src/core.cpp:417-456               -> k_kalman
//
// Reductions (J^T J, J^T F, score): RBV_REDUCE_FAST = per-thread FP64 partials + warp-shuffle tree in a fixed order;
// RBV_REDUCE_ORDERED = one lane per accumulator adds the float32 addends in keyline order, reproducing the CPU bits.
#include <stdlib.h>

#include "rbv_internal.cuh"

// ---------------------------------------------------------------------------------------------------------------
// One thread per keyline walks its normal line in |r| order (0, -1, 1, -2, 2, ...): the two divisions are done once per
// keyline, neighbouring threads are neighbouring keylines (coalesced SoA reads, nearby cells), and because every thread
// writes its nearest cells first, most far candidates are rejected by a plain load before they cost an atomic.
__global__ void __launch_bounds__(128) k_field_build(KeylineSoA K, const int* __restrict__ count, const float* __restrict__ threshold,
                                                     unsigned* __restrict__ field, const __grid_constant__ DevParams P) {
  const int b = blockIdx.y;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count[b]) return;
  const size_t o = (size_t)b * P.kmax + idx;
  const float gn = K.gnorm[o];
  const float thr = threshold[b];
  if (thr > 0.0f && gn < thr) return;                                                  // core.hpp:47
  const float2 g = K.grad[o], p = K.pos[o];
  const float ux = g.x / gn, uy = g.y / gn;
  unsigned* f = field + (size_t)b * P.npix;
  const unsigned lowkey = 0xFFFFFFu - (unsigned)idx;
  for (int j = 0; j < P.field_nr; ++j) {
    const int r = (j & 1) ? -((j + 1) >> 1) : (j >> 1);   // bijection onto [-a, a-1] (nr = 2a) or [-a, a] (nr = 2a+1)
    const int fi = rbv_pixel_index(P.rows, P.cols, uy * (float)r + p.y, ux * (float)r + p.x);  // :50-51
    if (fi < 0) continue;
    // sequential rule (:54-56): smallest |r| wins, ties go to the later (idx, r) -> min over (|r|, -idx)
    const unsigned key = ((unsigned)abs(r) << 24) | lowkey;
    if (__ldcg(f + fi) <= key) continue;         // cells only ever decrease: a stale read can only let a loser through to the atomic
    atomicMin(f + fi, key);
  }
}

// Band form of the same rule: one block owns `band_rows` image rows of one lane's field in SHARED memory. It scans the
// lane's keylines for those whose normal line can reach the band (|r| <= search range, so pos.y within reach + 1 rows),
// clips each line to the r-interval that lands in the band, and resolves "min (|r|, -idx)" with shared-memory atomics;
// the finished band is written to HBM once, coalesced (no separate reset pass, no global atomics). The result is the
// minimum over the same candidate set as k_field_build, hence identical for any keyline order.
#define FB_TPB 512
#define FB_G 8        // candidates per work item
#define FB_MAXI 12    // work items per keyline at most (search range <= 46)
__global__ void __launch_bounds__(FB_TPB) k_field_band(const float4* __restrict__ tgt, const float2* __restrict__ pos, const int* __restrict__ count,
                                                       const float* __restrict__ threshold, unsigned* __restrict__ field,
                                                       const __grid_constant__ DevParams P, int band_rows) {
  extern __shared__ __align__(16) unsigned fb_cell[];
  __shared__ int s_lo, s_hi;
  // per warp: the 32 keylines of its current round (unit normal, position, r-interval) and their work items
  __shared__ __align__(16) float4 s_kl[FB_TPB];       // (unit normal x, y, position x, y): one 16-byte load per work item
  __shared__ int s_ral[FB_TPB];                         // r-interval of the keyline: start (low 16 bits, signed) | length << 16
  __shared__ unsigned short s_item[FB_TPB * FB_MAXI];   // work item -> keyline of the warp's round (low 5 bits) | item of that keyline << 5
  const int b = blockIdx.y, t = threadIdx.x;
  const int rows = P.rows, cols = P.cols;
  const int r0 = blockIdx.x * band_rows, r1 = min(r0 + band_rows, rows);
  const int nrows = r1 - r0;
  const int ncell = nrows * cols;
  const int n = count[b];
  if ((ncell & 3) == 0) {
    uint4* c4 = reinterpret_cast<uint4*>(fb_cell);
    const uint4 e4 = make_uint4(RBV_FIELD_EMPTY, RBV_FIELD_EMPTY, RBV_FIELD_EMPTY, RBV_FIELD_EMPTY);
    for (int i = t; i < ncell / 4; i += FB_TPB) c4[i] = e4;
  } else {
    for (int i = t; i < ncell; i += FB_TPB) fb_cell[i] = RBV_FIELD_EMPTY;
  }
  if (t == 0) { s_lo = n; s_hi = 0; }
  __syncthreads();
  const int ra0 = P.field_r0, rb0 = P.field_r0 + P.field_nr - 1;
  const float reach = (float)max(-ra0, rb0) + 1.0f;
  const float ylo = (float)r0 - reach, yhi = (float)(r1 - 1) + reach;
  const size_t base = (size_t)b * P.kmax;
  {  // index range of the keylines that can reach the band (row-major ids make it a narrow range; any order is handled)
    int lo = n, hi = 0;
    for (int i0 = t; i0 < n; i0 += 4 * FB_TPB) {
      float y[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) y[u] = (i0 + u * FB_TPB < n) ? pos[base + i0 + u * FB_TPB].y : -1e30f;
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (y[u] >= ylo && y[u] <= yhi) { lo = min(lo, i0 + u * FB_TPB); hi = max(hi, i0 + u * FB_TPB + 1); }
    }
#pragma unroll
    for (int off = 16; off; off >>= 1) {
      lo = min(lo, __shfl_xor_sync(0xFFFFFFFFu, lo, off));
      hi = max(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, off));
    }
    if ((t & 31) == 0 && hi > lo) { atomicMin(&s_lo, lo); atomicMax(&s_hi, hi); }
  }
  __syncthreads();
  const int lo = s_lo, hi = s_hi;
  const float thr = threshold[b];
  const float r0f = (float)r0, r1f = (float)r1;
  // Every WARP works through rounds of 32 keylines on its own (no block barrier inside the loop: a warp that waits for its
  // keyline loads or its divisions does not hold the other 15 up - the block-wide rounds of the first version spent a third of
  // the kernel in barriers and load stalls). Phase 1 (lane = keyline): the unit normal - divided once per keyline by
  // k_pack_targets, the same float32 division - and the r-interval that can land in the band. Phase 2 (lane = work item of FB_G
  // consecutive r of one keyline): the intervals differ a lot in length (1 .. 80, by slope and by distance to the band), so they
  // are cut into equal items and dealt out densely over the warp's lanes.
  const int lane = t & 31, warp = t >> 5, wb = warp * 32;
  unsigned short* w_item = s_item + wb * FB_MAXI;
  for (int k0 = lo + wb; k0 < hi; k0 += FB_TPB) {
    const int idx = k0 + lane;
    int ra = 0, len = 0;
    float ux = 0.0f, uy = 0.0f;
    float2 p = make_float2(0.0f, 0.0f);
    if (idx < hi) {
      const float4 t0 = tgt[2 * (base + idx)], t1 = tgt[2 * (base + idx) + 1];           // (g.x, g.y, pos.x, pos.y), (g.x / |g|, g.y / |g|, |g|, -)
      p = make_float2(t0.z, t0.w);
      const float gn = t1.z;
      if (p.y >= ylo && p.y <= yhi && !(thr > 0.0f && gn < thr)) {                     // core.hpp:47
        ux = t1.x; uy = t1.y;
        int rb = rb0;
        ra = ra0;
        if (fabsf(uy) > 1e-2f) {   // rows hit are monotone in r: clip to the r-interval that can land in [r0, r1) (+-1 margin, re-tested below;
          const float iuy = __fdividef(1.0f, uy);   // an approximate reciprocal moves the bounds by < 1e-2, far inside the margin)
          const float t0c = (r0f - 0.5f - p.y) * iuy, t1c = (r1f - 0.5f - p.y) * iuy;
          ra = max(ra, __float2int_rd(fminf(t0c, t1c)) - 1);
          rb = min(rb, __float2int_ru(fmaxf(t0c, t1c)) + 1);
        }
        len = max(rb - ra + 1, 0);
        if (!(fabsf(ux) <= 2.0f && fabsf(uy) <= 2.0f)) len = 0;   // non-finite normal: the reference's getIndex yields -1 for every r
      }
    }
    s_kl[t] = make_float4(ux, uy, p.x, p.y); s_ral[t] = (ra & 0xFFFF) | (len << 16);
    // exclusive scan of the item counts over the warp, then every keyline lists its items in the warp's item table
    const int items = min((len + FB_G - 1) / FB_G, FB_MAXI);
    int inc = items;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int v = __shfl_up_sync(0xFFFFFFFFu, inc, off);
      if (lane >= off) inc += v;
    }
    const int first = inc - items;
    const int total = __shfl_sync(0xFFFFFFFFu, inc, 31);
    for (int i = 0; i < items; ++i) w_item[first + i] = (unsigned short)(lane | (i << 5));
    __syncwarp();
    for (int j = lane; j < total; j += 32) {
      const int it = w_item[j];
      const int ks = wb + (it & 31);
      const int ral = s_ral[ks];
      const int kra = (int)(short)(ral & 0xFFFF);
      const int rs = kra + (it >> 5) * FB_G;
      const int re = kra + (ral >> 16);
      const float4 kl = s_kl[ks];
      const float kux = kl.x, kuy = kl.y, kpx = kl.z, kpy = kl.w;
      const unsigned lowkey = 0xFFFFFFu - (unsigned)(k0 + (it & 31));
      // The FB_G candidates of the item as straight-line predicated code: FB_G independent address chains, then FB_G
      // loads, then the (few) atomics. A loop with early `continue`s made every warp crawl through one dependent chain
      // with a branch bubble every few instructions (measured ~12 cycles per issued instruction).
      // getIndex (core.hpp:66-71): (int)std::round(v), half away from zero = sign(v) * trunc(|v| + 0.5). All on the FP32
      // add pipe: the RZ add of 0.5 is what roundf() itself does, the RZ add of 2^23 leaves trunc() in the mantissa
      // (values >= 2^23, Inf and NaN come out as huge indices and fail the range test). No FRND / F2I conversions.
      // The range tests are made on the float coordinate, not on the rounded index (round-half-away(v) in [a, b) <=> v in [a - 0.5, b - 0.5),
      // and v > -0.5 for a = 0): four chained float compares per candidate and no integer index at all - the cell address comes straight
      // from the bit patterns (bits = 0x4B000000 + index): addr = bits_y * 4 cols + bits_x * 4 + A0 with the constants folded into A0
      // (unsigned wrap-around arithmetic). |r| of the key the same way: bits(|fr| + 2^23) << 24 drops the 0x4B.
      const float fr0 = (float)rs, fre = (float)re;
      const unsigned cell_a = (unsigned)__cvta_generic_to_shared(fb_cell);
      const unsigned cols4 = 4u * (unsigned)cols;
      const unsigned A0 = cell_a - cols4 * (0x4B000000u + (unsigned)r0) - 4u * 0x4B000000u;
      const float ylo = r0 > 0 ? __int_as_float(__float_as_int(r0f - 0.5f) - 1) : -0.5f;   // fy > ylo  <=>  round(fy) >= r0 (r0 - 0.5 itself rounds to r0)
      const float yhi = r1f - 0.5f, xhi = (float)cols - 0.5f;                                // fy < yhi  <=>  round(fy) < r1
      unsigned addr[FB_G], key[FB_G], cur[FB_G];
#pragma unroll
      for (int u = 0; u < FB_G; ++u) {
        float fr;   // (float)(rs + u), exact; as one FADD (left to itself nvcc converts rs + u again for every candidate: IADD + I2F)
        asm("add.rn.f32 %0, %1, %2;" : "=f"(fr) : "f"(fr0), "f"((float)u));
        const float fy = kuy * fr + kpy;                                                  // :50
        const float fx = kux * fr + kpx;                                                  // :51
        const unsigned by = (unsigned)__float_as_int(__fadd_rz(__fadd_rz(fabsf(fy), 0.5f), 8388608.0f));
        const unsigned bx = (unsigned)__float_as_int(__fadd_rz(__fadd_rz(fabsf(fx), 0.5f), 8388608.0f));
        const unsigned kb = (unsigned)__float_as_int(fabsf(fr) + 8388608.0f);
        // address and key as explicit multiply-adds (left to the compiler the folded constants come apart again: 7 integer instructions
        // per candidate instead of 3): addr = by * 4 cols + (bx * 4 + A0); key = |r| << 24 | (0xFFFFFF - keyline), :54-56 as min over (|r|, -idx)
        asm("{ .reg .u32 t;\n"
            "  mad.lo.u32 t, %3, 4, %5;\n"
            "  mad.lo.u32 %0, %2, %4, t;\n"
            "  mad.lo.u32 %1, %6, 16777216, %7; }"
            : "=r"(addr[u]), "=r"(key[u]) : "r"(by), "r"(bx), "r"(cols4), "r"(A0), "r"(kb), "r"(lowkey));
        // in range: r < re, round(fy) in [r0, r1), round(fx) in [0, cols) - a negative coordinate rounds to an index >= 0 only for
        // v > -0.5 (round(-0.3) = -0 -> 0, round(-0.5) = -1); NaN fails every compare. The five compares chain into ONE predicate that
        // guards the load (a predicated instruction, no branch per candidate); out of range reads as 0 = "nothing to improve" below.
        asm volatile("{ .reg .pred p;\n"
                     "  setp.lt.f32 p, %2, %3;\n"
                     "  setp.gt.and.f32 p, %4, %5, p;\n"
                     "  setp.lt.and.f32 p, %4, %6, p;\n"
                     "  setp.gt.and.f32 p, %7, 0fBF000000, p;\n"
                     "  setp.lt.and.f32 p, %7, %8, p;\n"
                     "  mov.u32 %0, 0;\n"
                     "  @p ld.volatile.shared.u32 %0, [%1]; }"
                     : "=r"(cur[u]) : "r"(addr[u]), "f"(fr), "f"(fre), "f"(fy), "f"(ylo), "f"(yhi), "f"(fx), "f"(xhi));
      }
      // compares + atomics, all loads issued before the first use
#pragma unroll
      for (int u = 0; u < FB_G; ++u)
        asm volatile("{ .reg .pred q;\n"
                     "  setp.gt.u32 q, %2, %1;\n"
                     "  @q red.shared.min.u32 [%0], %1; }" ::"r"(addr[u]), "r"(key[u]), "r"(cur[u]) : "memory");
    }
    __syncwarp();   // the warp's tables are rewritten by its next round
  }
  __syncthreads();
  unsigned* dst = field + (size_t)b * P.npix + (size_t)r0 * cols;
  if ((cols & 3) == 0) {
    const uint4* s4 = reinterpret_cast<const uint4*>(fb_cell);
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    for (int i = t; i < ncell / 4; i += FB_TPB) d4[i] = s4[i];
  } else {
    for (int i = t; i < ncell; i += FB_TPB) dst[i] = fb_cell[i];
  }
}

// rows per band: as many as fit ~100 KB (two blocks per SM); below 16 rows the keyline scan would be repeated too often
static int field_band_rows(const rbv_ctx* c) {
  const int by_smem = (90 * 1024) / (c->P.cols * 4);   // + ~23 KB of static tables: two blocks fit the 227 KB of an SM
  int h = by_smem > 32 ? 32 : by_smem;
  if (c->tune[RBV_TUNE_FIELD_ROWS] > 0) return c->tune[RBV_TUNE_FIELD_ROWS] < h ? c->tune[RBV_TUNE_FIELD_ROWS] : h;
  if (h >= 16) {   // prefer a row count that divides the image: no short last band
    for (int d = h; d >= h - 4 && d >= 16; --d)
      if (c->P.rows % d == 0) { h = d; break; }
  }
  // few lanes: the grid (bands x lanes) should still cover the 148 SMs; lower bands repeat the keyline scan more often but
  // a single stream is bound by the longest block, not by total work
  while (h > 8 && ((c->P.rows + h - 1) / h) * c->P.lanes < 148) h = (h + 1) / 2;
  return h;
}

// What tryVel reads of a matched target keyline (core.cpp:39-44,62-75), packed into one 32-byte sector per keyline so
// that the gather behind the distance-field lookup is one memory transaction; the unit gradient is divided once here
// instead of once per keyline per LM iteration (same float division, same result).
__global__ void k_pack_targets(KeylineSoA K, const int* __restrict__ count, float4* __restrict__ tgt, int kmax) {
  const int b = blockIdx.y, idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count[b]) return;
  const size_t o = (size_t)b * kmax + idx;
  const float2 g = K.grad[o], p = K.pos[o];
  const float gn = K.gnorm[o];
  tgt[2 * o] = make_float4(g.x, g.y, p.x, p.y);
  tgt[2 * o + 1] = make_float4(g.x / gn, g.y / gn, gn, 0.0f);                            // core.cpp:64-65 (and |g| for k_field_band)
}

int tr_pack_targets(rbv_ctx* c, rbv_map* map) {
  const DevParams& P = c->P;
  RBV_LAUNCH(c, k_pack_targets, (dim3((P.keylines_max + 255) / 256, P.lanes)), (256), 0, map->k, map->count, c->tgt, P.kmax);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

int tr_init_ordered(void);
int tr_init(rbv_ctx* c) {
  (void)c;   // the current device is the context's
  RBV_CUDA(cudaFuncSetAttribute(k_field_band, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  return tr_init_ordered();
}

int tr_build_field(rbv_ctx* c, rbv_map* map, bool packed) {
  const DevParams& P = c->P;
  const int H = field_band_rows(c);
  const bool band = (c->tune[RBV_TUNE_FIELD] == 2 || (c->tune[RBV_TUNE_FIELD] == 0 && H >= 8)) && P.field_nr <= FB_G * FB_MAXI;
  if (!packed) { int rc = tr_pack_targets(c, map); if (rc != RBV_OK) return rc; }   // first: the band kernel reads the packed records (unit normals divided once); packed: k_detect_emit has written them
  if (band && H >= 1) {
    const size_t smem = (size_t)H * P.cols * 4;
    RBV_LAUNCH(c, k_field_band, (dim3((P.rows + H - 1) / H, P.lanes)), (FB_TPB), smem, c->tgt, map->k.pos, map->count, map->threshold, c->field, P, H);
  } else {
    { int rc = rbv_fill32(c, c->field, RBV_FIELD_EMPTY, (size_t)P.lanes * P.npix); if (rc != RBV_OK) return rc; }
    RBV_LAUNCH(c, k_field_build, (dim3((P.keylines_max + 127) / 128, P.lanes)), (128), 0, map->k, map->count, map->threshold, c->field, P);
  }
  c->field_map = map;
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// rotateKeylines: R * makeVector(float, float, double) is a float x double product accumulated in double
// (TooN make_vector.hh:26, operators.hh:200-208) and rounded to float on assignment.
__global__ void k_rotate(KeylineSoA K, const int* __restrict__ count, const float* __restrict__ lane_in, int* zero_cnt,
                         unsigned long long* zero_probes, const int* __restrict__ nkey_src, const float* __restrict__ athr_src,
                         const __grid_constant__ DevParams P) {
  const int b = blockIdx.y, idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (zero_cnt && idx == 0) {   // first kernel of a fused tracker sequence: reset the lane's match counters for the stages behind it
    zero_cnt[b * CNT_STRIDE + CNT_FWD] = 0; zero_cnt[b * CNT_STRIDE + CNT_DIR] = 0;
    zero_cnt[b * CNT_STRIDE + CNT_KF] = 0; zero_cnt[b * CNT_STRIDE + CNT_REG] = 0;
    zero_probes[b] = 0ull;
    if (nkey_src) {   // the new map's keyline count and auto threshold travel with the counters (one D2H for the frame)
      zero_cnt[b * CNT_STRIDE + CNT_NKEY] = nkey_src[b];
      zero_cnt[b * CNT_STRIDE + CNT_ATHR] = __float_as_int(athr_src[b]);
    }
  }
  if (idx >= count[b]) return;
  const float* R = lane_in + (size_t)b * IN_STRIDE + IN_R;
  const size_t o = (size_t)b * P.kmax + idx;
  const float fm = P.fm;
  float2 pi = K.pos_img[o];
  const float rho = K.rho[o], srho = K.srho[o];   // every load of the keyline first: one memory round trip (the stores below may alias for all the compiler knows)
  const float2 g = K.grad[o];
  const double v0 = (double)(pi.x / fm), v1 = (double)(pi.y / fm), v2 = 1.0;
  float q[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    double acc = 0.0;
    acc += (double)R[r * 3 + 0] * v0;
    acc += (double)R[r * 3 + 1] * v1;
    acc += (double)R[r * 3 + 2] * v2;
    q[r] = (float)acc;
  }
  if (fabsf(q[2]) > 0.0f) {
    pi.x = q[0] / q[2] * fm;
    pi.y = q[1] / q[2] * fm;
    K.pos_img[o] = pi;
    K.rho[o] = rho / q[2];
    K.srho[o] = srho / q[2];
  }
  float2 go;
  {
    double acc = 0.0;
    acc += (double)R[0] * (double)g.x; acc += (double)R[1] * (double)g.y; acc += (double)R[2] * 0.0;
    go.x = (float)acc;
    acc = 0.0;
    acc += (double)R[3] * (double)g.x; acc += (double)R[4] * (double)g.y; acc += (double)R[5] * 0.0;
    go.y = (float)acc;
  }
  K.grad[o] = go;
}

int tr_rotate(rbv_ctx* c, rbv_map* map) {
  const DevParams& P = c->P;
  RBV_LAUNCH(c, k_rotate, (dim3((P.keylines_max + 255) / 256, P.lanes)), (256), 0, map->k, map->count, c->lane_in,
             c->seq_fused ? c->lane_cnt : nullptr, c->probes, (c->seq_fused && c->seq_new) ? c->seq_new->count : nullptr,
             (c->seq_fused && c->seq_new) ? c->seq_new->threshold : nullptr, P);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024) k_quantile(KeylineSoA K, const int* __restrict__ count, float percentile, int bins, int kmax,
                                                  LMState* lm, float* lane_out, int to_lm, const float* __restrict__ lane_in, int* lm_done) {
  __shared__ int hist[RBV_MAX_QBINS];
  const int b = blockIdx.x, n = count[b];
  for (int i = threadIdx.x; i < bins; i += blockDim.x) hist[i] = 0;
  __syncthreads();
  const float span = RBV_RHO_MAX - RBV_RHO_MIN;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    int q = rbv_f2i((float)bins * (K.srho[(size_t)b * kmax + i] - RBV_RHO_MIN) / span);   // edge_map.cpp:42
    q = q > bins - 1 ? bins - 1 : q;
    q = q < 0 ? 0 : q;
    atomicAdd(&hist[q], 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float out = 1e3f;
    const float lim = percentile * (float)n;
    for (int i = 0, a = 0; i < bins; ++i) {
      if ((float)a > lim) { out = (float)i * span / (float)bins + RBV_RHO_MIN; break; }    // :49-50
      a += hist[i];
    }
    if (to_lm) {
      lm[b].sigma_rho_min = out;
      lm_done[b] = 0;   // k_minimize_ord*: number of finished LM calls of this lane (the minimisation always follows its quantile)
      if (to_lm == 2)   // also the start velocity of Core::minimizeVel (what k_lm_init does for the stand-alone call)
        for (int i = 0; i < 3; ++i) lm[b].vel[i] = lane_in[(size_t)b * IN_STRIDE + IN_VEL + i];
    }
    lane_out[(size_t)b * OUT_STRIDE + OUT_Q] = out;
  }
}

int tr_quantile(rbv_ctx* c, rbv_map* map, float pct, int bins, int to_lm) {
  if (bins < 1 || bins > RBV_MAX_QBINS) { rbv_set_error("quantile bins out of range"); return RBV_ERR_ARG; }
  RBV_LAUNCH(c, k_quantile, (c->P.lanes), (1024), 0, map->k, map->count, pct, bins, c->P.kmax, c->lm, c->lane_out, to_lm, c->lane_in, c->lm_done);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// tryVel, per-keyline part (core.cpp:86-141). Addends (zero for skipped keylines: x + 0 == x exactly):
//   0 score f*f | 1..6 JtJ 00,11,22,01,02,12 | 7..9 JtF 0,1,2
// code for the residual rule: >= 0 matched (|fi|), -1 in-image but unmatched, -2 no residual write.
struct TryVelIn {
  float v0, v1, v2, sigma_rho_min, thr;
};
__device__ __forceinline__ void tryvel_point(const KeylineSoA& Ko, const float4* __restrict__ tgt, const unsigned* __restrict__ field_b,
                                             size_t o, size_t base, float res, const TryVelIn& in, const DevParams& P,
                                             float* add, float& code, int& fwd) {
#pragma unroll
  for (int a = 0; a < 10; ++a) add[a] = 0.0f;
  code = -2.0f;
  fwd = -1;
  const float v0 = in.v0, v1 = in.v1, v2 = in.v2;
  const float gn = Ko.gnorm[o];
  const float srho = Ko.srho[o];
  const bool skip = (in.thr > 0.0f && gn < in.thr) || (srho > in.sigma_rho_min) || (Ko.matches[o] < P.min_matches);   // core.cpp:88,91
  if (skip) return;
  float w = 1.0f;
  if (res > P.reweight_distance) w = P.reweight_distance / res;                        // :94
  const float rho = Ko.rho[o];
  const float z_p = (float)(1.0 / (double)rho + (double)v2);                           // :96
  const float2 pi = Ko.pos_img[o];
  bool penalty = z_p <= 0.0f;
  float rho_p = 0.0f, px = 0.0f, py = 0.0f, pxc = 0.0f, pyc = 0.0f;
  int x = 0, y = 0;
  if (!penalty) {
    rho_p = (float)(1.0 / (double)z_p);                                                // :104
    px = rho_p * (v0 * P.fm - v2 * pi.x) + pi.x;                                       // :105
    py = rho_p * (v1 * P.fm - v2 * pi.y) + pi.y;
    pxc = px + P.cx; pyc = py + P.cy;
    x = rbv_d2i((double)pxc + 0.5);                                                    // :111
    y = rbv_d2i((double)pyc + 0.5);
    penalty = (x < 1 || y < 1 || x >= P.cols - 1 || y >= P.rows - 1);                  // :114
  }
  if (penalty) {
    const float f = (float)((1.0 / (double)srho) * (double)P.search_range * (double)w);   // :99 / :115
    add[0] = f * f;
    return;
  }
  float dfx = 0.0f, dfy = 0.0f, f;
  bool matched = false;
  const unsigned cell = field_b[(size_t)y * P.cols + x];
  if (cell != RBV_FIELD_EMPTY) {
    const int tid = (int)(0xFFFFFFu - (cell & 0xFFFFFFu));
    const float4 t0 = tgt[2 * (base + tid)];                                           // (grad.x, grad.y, pos.x, pos.y) of the target
    const float2 tg = make_float2(t0.x, t0.y);
    const float2 g = Ko.grad[o];
    const float n2 = gn * gn;                                                          // testfk :40-42
    const float dp = tg.x * g.x + tg.y * g.y;
    if (!(fabsf(dp - n2) > P.match_treshold * n2)) {
      matched = true;
      const float4 t1 = tgt[2 * (base + tid) + 1];                                     // (grad / gradient_norm, -, -): same 32-byte sector
      const float2 tp = make_float2(t0.z, t0.w);
      const float dx = pxc - tp.x, dy = pyc - tp.y;                                    // :62-63
      const float gnx = t1.x, gny = t1.y;                                              // tg / tgn, divided once when the field is built
      const float fi = dx * gnx + dy * gny;                                            // :67
      dfx = gnx / srho; dfy = gny / srho;
      fwd = tid;                                                                       // :73
      f = fi / srho;                                                                   // :75
      code = fabsf(fi);                                                                // :141
    }
  }
  if (!matched) { f = P.search_range / srho; code = -1.0f; }                           // :52 / :59
  f *= w;
  add[0] = f * f;
  const float jx = rho_p * P.fm * dfx * w;                                             // :126-128
  const float jy = rho_p * P.fm * dfy * w;
  const float jz = -rho_p * (px * dfx + py * dfy) * w;
  add[1] = jx * jx; add[2] = jy * jy; add[3] = jz * jz;
  add[4] = jx * jy; add[5] = jx * jz; add[6] = jy * jz;
  add[7] = jx * f; add[8] = jy * f; add[9] = jz * f;
}

// ORDERED mode, phase 1: addends to prod[10][B][kmax], residual code to fiabs
__global__ void __launch_bounds__(256) k_tryvel_eval(KeylineSoA Ko, const int* __restrict__ count_o, const float* __restrict__ thr_o,
                                                     const float4* __restrict__ tgt, const unsigned* __restrict__ field, const float* __restrict__ vel_ptr,
                                                     int vel_stride, const float* __restrict__ sr_ptr, int sr_stride,
                                                     float* __restrict__ prod, float* __restrict__ fiabs,
                                                     const __grid_constant__ DevParams P) {
  const int b = blockIdx.y, idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count_o[b]) return;
  const size_t base = (size_t)b * P.kmax, o = base + idx;
  const size_t ps = (size_t)P.lanes * P.kmax;
  TryVelIn in;
  in.v0 = vel_ptr[(size_t)b * vel_stride]; in.v1 = vel_ptr[(size_t)b * vel_stride + 1]; in.v2 = vel_ptr[(size_t)b * vel_stride + 2];
  in.sigma_rho_min = sr_ptr[(size_t)b * sr_stride];
  in.thr = thr_o[b];
  float add[10], code;
  int fwd;
  tryvel_point(Ko, tgt, field + (size_t)b * P.npix, o, base, Ko.residual[o], in, P, add, code, fwd);
  Ko.match_fwd[o] = fwd;                                                                 // :87
  fiabs[o] = code;
#pragma unroll
  for (int a = 0; a < 10; ++a) prod[a * ps + o] = add[a];
}

// block-wide sum of NV doubles per thread in a fixed order (deterministic): shuffle tree inside a warp, then the
// first NV threads add the 32 warp partials in warp order.
template <int NV>
__device__ __forceinline__ void block_sum_fixed(double* acc, double* sm /*[32][NV]*/, double* result /*[NV] in smem*/) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int a = 0; a < NV; ++a) {
    double v = acc[a];
#pragma unroll
    for (int off = 16; off; off >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, off);
    if (lane == 0) sm[warp * NV + a] = v;
  }
  __syncthreads();
  if (threadIdx.x < NV) {
    double s = 0.0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += sm[w * NV + threadIdx.x];
    result[threadIdx.x] = s;
  }
  __syncthreads();
}

// LM bookkeeping of Core::minimizeVel (core.cpp:159-185), executed by one thread per lane after each tryVel
__device__ void lm_update(LMState& S, int call, int iterations, float score, const float* JtJn, const float* JtFn, const float* vel_used) {
  float* tr = S.trace + call * 16;
  for (int i = 0; i < 3; ++i) tr[i] = vel_used[i];
  for (int i = 0; i < 9; ++i) tr[3 + i] = JtJn[i];
  for (int i = 0; i < 3; ++i) tr[12 + i] = JtFn[i];
  tr[15] = score;
  bool need_step;
  if (call == 0) {
    S.F = score;
    for (int i = 0; i < 9; ++i) S.JtJ[i] = JtJn[i];
    for (int i = 0; i < 3; ++i) S.JtF[i] = JtFn[i];
    S.v = 2.0f;
    float mx = JtJn[0];
    for (int i = 1; i < 9; ++i) if (JtJn[i] > mx) mx = JtJn[i];                          // TooN::max_element
    S.u = 1e-3f * mx;                                                                    // :161-163
    need_step = iterations > 0;
  } else {
    double den = 0.0;                                                                    // 0.5*h * (u*h - JtF), :172
    for (int i = 0; i < 3; ++i) {
      const float d = S.u * S.h[i] - S.JtF[i];
      den += (0.5 * (double)S.h[i]) * (double)d;
    }
    const float gain = (float)((double)(S.F - score) / den);
    if (gain > 0.0f) {                                                                   // :173-179
      S.F = score;
      for (int i = 0; i < 3; ++i) S.vel[i] = S.vnew[i];
      for (int i = 0; i < 9; ++i) S.JtJ[i] = JtJn[i];
      for (int i = 0; i < 3; ++i) S.JtF[i] = JtFn[i];
      const double t = 2.0 * (double)gain - 1.0;
      const double s = 1.0 - (t * t * t);
      S.u = (float)((double)S.u * (0.33 > s ? 0.33 : s));
      S.v = 2.0f;
    } else {                                                                             // :180-183
      S.u *= S.v;
      S.v = (float)((double)S.v * 2.0);
    }
    need_step = call < iterations;
  }
  if (need_step) {
    float ApI[9], inv[9];
    for (int i = 0; i < 9; ++i) ApI[i] = S.JtJ[i];
    ApI[0] = S.JtJ[0] + S.u; ApI[4] = S.JtJ[4] + S.u; ApI[8] = S.JtJ[8] + S.u;           // :167
    rbv_inv3(ApI, inv);
    const float g[3] = {S.JtF[0] * -1.0f, S.JtF[1] * -1.0f, S.JtF[2] * -1.0f};
    for (int r = 0; r < 3; ++r) {                                                        // :168
      float acc = 0.0f;
      for (int j = 0; j < 3; ++j) acc += inv[r * 3 + j] * g[j];
      S.h[r] = acc;
    }
    for (int i = 0; i < 3; ++i) S.vnew[i] = S.vel[i] + S.h[i];                           // :169
  } else {
    rbv_inv3(S.JtJ, S.Rvel);                                                             // :185
  }
}

// tryVel, reduction part: one block per lane. (1) residual rule incl. the reference's carried `fi` (see DESIGN.md:
// an in-image unmatched keyline stores |fi| of the most recent matched keyline, 0 before the first match);
// (2) the ten sums; (3) LM step or plain output.
__global__ void __launch_bounds__(1024) k_tryvel_reduce(KeylineSoA Ko, const int* __restrict__ count_o, const float* __restrict__ prod,
                                                        const float* __restrict__ fiabs, LMState* lm, float* lane_out,
                                                        const float* __restrict__ vel_ptr, int vel_stride, int lm_call, int ordered,
                                                        const __grid_constant__ DevParams P) {
  __shared__ double sm[32 * 10];
  __shared__ double result[10];
  __shared__ float warp_val[32];
  __shared__ int warp_has[32];
  __shared__ float carry_val;
  __shared__ int carry_has;
  const int b = blockIdx.x, n = count_o[b], t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const size_t base = (size_t)b * P.kmax;
  const size_t ps = (size_t)P.lanes * P.kmax;
  if (t == 0) { carry_val = 0.0f; carry_has = 0; }
  __syncthreads();
  // (1) "last matched |fi|" propagation in keyline order
  for (int i0 = 0; i0 < n; i0 += 1024) {
    const int idx = i0 + t;
    const float code = (idx < n) ? fiabs[base + idx] : -2.0f;
    int has = code >= 0.0f;
    float val = has ? code : 0.0f;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int h2 = __shfl_up_sync(0xFFFFFFFFu, has, off);
      const float v2 = __shfl_up_sync(0xFFFFFFFFu, val, off);
      if (lane >= off && !has) { has = h2; val = v2; }
    }
    if (lane == 31) { warp_has[warp] = has; warp_val[warp] = val; }
    __syncthreads();
    if (!has) {
      for (int w = warp - 1; w >= 0; --w)
        if (warp_has[w]) { has = 1; val = warp_val[w]; break; }
      if (!has && carry_has) { has = 1; val = carry_val; }
    }
    if (idx < n && code >= -1.5f) Ko.residual[base + idx] = has ? val : 0.0f;            // core.cpp:141
    __syncthreads();
    if (t == 1023) { if (has) { carry_has = 1; carry_val = val; } }
    __syncthreads();
  }
  // (2) sums
  if (ordered) {
    if (warp == 0 && lane < 10) {   // one lane per accumulator: same instruction stream, ten independent serial chains
      const float* p = prod + lane * ps + base;
      float s = 0.0f;
      int i = 0;
      for (; i + 8 <= n; i += 8) {
        const float4 a = *reinterpret_cast<const float4*>(p + i);
        const float4 c = *reinterpret_cast<const float4*>(p + i + 4);
        s += a.x; s += a.y; s += a.z; s += a.w; s += c.x; s += c.y; s += c.z; s += c.w;
      }
      for (; i < n; ++i) s += p[i];
      result[lane] = (double)s;
    }
    __syncthreads();
  } else {
    double acc[10];
#pragma unroll
    for (int a = 0; a < 10; ++a) acc[a] = 0.0;
    for (int i = t; i < n; i += 1024) {
#pragma unroll
      for (int a = 0; a < 10; ++a) acc[a] += (double)prod[a * ps + base + i];
    }
    block_sum_fixed<10>(acc, sm, result);
  }
  // (3)
  if (t == 0) {
    float JtJ[9], JtF[3];
    const float score = (float)result[0];
    JtJ[0] = (float)result[1]; JtJ[4] = (float)result[2]; JtJ[8] = (float)result[3];
    JtJ[1] = JtJ[3] = (float)result[4]; JtJ[2] = JtJ[6] = (float)result[5]; JtJ[5] = JtJ[7] = (float)result[6];
    JtF[0] = (float)result[7]; JtF[1] = (float)result[8]; JtF[2] = (float)result[9];
    if (lm_call >= 0) {
      float vu[3] = {vel_ptr[(size_t)b * vel_stride], vel_ptr[(size_t)b * vel_stride + 1], vel_ptr[(size_t)b * vel_stride + 2]};
      lm_update(lm[b], lm_call, P.iterations, score, JtJ, JtF, vu);
    }
    float* out = lane_out + (size_t)b * OUT_STRIDE;
    for (int i = 0; i < 9; ++i) out[OUT_JTJ + i] = JtJ[i];
    for (int i = 0; i < 3; ++i) out[OUT_JTF + i] = JtF[i];
    out[OUT_SCORE] = score;
  }
}

// ---- RBV_REDUCE_FAST: tryVel in ONE kernel -----------------------------------------------------------------------------
// grid (blocks of 256 keylines, lanes). Each block evaluates its keylines, reduces the ten addends in FP64 (shuffle tree
// + fixed warp order) into partials[lane][block][10]; the block that takes the last ticket of its lane adds the block
// partials in block order (a fixed order -> deterministic) and runs the LM step. The reference's carried residual
// ("an in-image unmatched keyline stores |fi| of the most recent matched keyline") is resolved inside the block; a
// keyline with no earlier match in its block stores the marker -1 and looks back over the previous blocks' "last
// matched |fi|" (blkagg, ping-pong per call) when the next call reads it. k_residual_resolve clears the markers after
// the last call so the map never leaves the library with one.
#define TV_TPB 256
#ifndef TV_MINB
#define TV_MINB 1   // minimum resident blocks per SM asked of ptxas for the tryVel evaluation kernels (A/B knob)
#endif
__device__ __forceinline__ float residual_lookback(const float* __restrict__ agg_lane, int blk) {
  for (int j = blk - 1; j >= 0; --j) {
    const float a = __ldcg(agg_lane + j);
    if (a >= 0.0f) return a;
  }
  return 0.0f;
}

__global__ void __launch_bounds__(TV_TPB) k_tryvel_fast(KeylineSoA Ko, const int* __restrict__ count_o, const float* __restrict__ thr_o,
                                                        const float4* __restrict__ tgt, const unsigned* __restrict__ field, const float* __restrict__ vel_ptr,
                                                        int vel_stride, const float* __restrict__ sr_ptr, int sr_stride,
                                                        double* __restrict__ partials, unsigned* __restrict__ ticket,
                                                        const float* __restrict__ agg_prev, float* __restrict__ agg_cur, int nblk,
                                                        LMState* lm, float* lane_out, int lm_call, const __grid_constant__ DevParams P) {
  __shared__ double sm[(TV_TPB / 32) * 10];
  __shared__ double result[10];
  __shared__ float warp_val[TV_TPB / 32];
  __shared__ int warp_has[TV_TPB / 32];
  __shared__ int s_last;
  const int b = blockIdx.y, n = count_o[b], t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int nact = max(1, (n + TV_TPB - 1) / TV_TPB);
  if ((int)blockIdx.x >= nact) return;
  const int idx = blockIdx.x * TV_TPB + t;
  const size_t base = (size_t)b * P.kmax, o = base + idx;
  float add[10], code = -2.0f, res = 0.0f;
  bool was_pending = false;
#pragma unroll
  for (int a = 0; a < 10; ++a) add[a] = 0.0f;
  if (idx < n) {
    TryVelIn in;
    in.v0 = vel_ptr[(size_t)b * vel_stride]; in.v1 = vel_ptr[(size_t)b * vel_stride + 1]; in.v2 = vel_ptr[(size_t)b * vel_stride + 2];
    in.sigma_rho_min = sr_ptr[(size_t)b * sr_stride];
    in.thr = thr_o[b];
    res = Ko.residual[o];
    if (res < 0.0f) { was_pending = true; res = residual_lookback(agg_prev + (size_t)b * nblk, blockIdx.x); }
    int fwd;
    tryvel_point(Ko, tgt, field + (size_t)b * P.npix, o, base, res, in, P, add, code, fwd);
    Ko.match_fwd[o] = fwd;                                                               // core.cpp:87
  }
  // carried |fi| inside the block, keyline order
  int has = code >= 0.0f;
  float val = has ? code : 0.0f;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int h2 = __shfl_up_sync(0xFFFFFFFFu, has, off);
    const float v2 = __shfl_up_sync(0xFFFFFFFFu, val, off);
    if (lane >= off && !has) { has = h2; val = v2; }
  }
  if (lane == 31) { warp_has[warp] = has; warp_val[warp] = val; }
  // ten sums: shuffle tree per warp
#pragma unroll
  for (int a = 0; a < 10; ++a) {
    double v = (double)add[a];
#pragma unroll
    for (int off = 16; off; off >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, off);
    if (lane == 0) sm[warp * 10 + a] = v;
  }
  __syncthreads();
  if (!has) {
    for (int w = warp - 1; w >= 0; --w)
      if (warp_has[w]) { has = 1; val = warp_val[w]; break; }
  }
  if (idx < n) {
    if (code >= -1.5f) Ko.residual[o] = has ? val : (blockIdx.x == 0 ? 0.0f : -1.0f);    // core.cpp:141 (-1 = look back next time)
    else if (was_pending) Ko.residual[o] = res;                                          // no write this call: keep the resolved value
  }
  if (t == TV_TPB - 1) agg_cur[(size_t)b * nblk + blockIdx.x] = has ? val : -1.0f;
  if (t < 10) {
    double s2 = 0.0;
#pragma unroll
    for (int w = 0; w < TV_TPB / 32; ++w) s2 += sm[w * 10 + t];
    partials[((size_t)b * nblk + blockIdx.x) * 10 + t] = s2;
  }
  __threadfence();
  __syncthreads();
  if (t == 0) s_last = (atomicAdd(&ticket[b], 1u) == (unsigned)(nact - 1));
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (t < 10) {
    double s2 = 0.0;
    for (int j = 0; j < nact; ++j) s2 += __ldcg(partials + ((size_t)b * nblk + j) * 10 + t);
    result[t] = s2;
  }
  __syncthreads();
  if (t == 0) {
    ticket[b] = 0u;
    float JtJ[9], JtF[3];
    const float score = (float)result[0];
    JtJ[0] = (float)result[1]; JtJ[4] = (float)result[2]; JtJ[8] = (float)result[3];
    JtJ[1] = JtJ[3] = (float)result[4]; JtJ[2] = JtJ[6] = (float)result[5]; JtJ[5] = JtJ[7] = (float)result[6];
    JtF[0] = (float)result[7]; JtF[1] = (float)result[8]; JtF[2] = (float)result[9];
    if (lm_call >= 0) {
      float vu[3] = {vel_ptr[(size_t)b * vel_stride], vel_ptr[(size_t)b * vel_stride + 1], vel_ptr[(size_t)b * vel_stride + 2]};
      lm_update(lm[b], lm_call, P.iterations, score, JtJ, JtF, vu);
    }
    float* out = lane_out + (size_t)b * OUT_STRIDE;
    for (int i = 0; i < 9; ++i) out[OUT_JTJ + i] = JtJ[i];
    for (int i = 0; i < 3; ++i) out[OUT_JTF + i] = JtF[i];
    out[OUT_SCORE] = score;
  }
}

// ---- the same, KPT consecutive keylines per thread ---------------------------------------------------------------------
// k_tryvel_fast is bound by the latency of a dependent chain (keyline fields -> field cell -> target record) whose loads sit
// behind the reference's early-outs. Here every thread owns KPT consecutive keylines (thread order = keyline order, so the
// carried-residual rule and the fixed summation order survive), fetches all their fields with vector loads BEFORE any test,
// then issues the KPT field-cell gathers together and the KPT target gathers together: three memory round trips per thread
// with KPT requests in flight in each, instead of five with one. The per-keyline arithmetic is tryvel_point's, statement by
// statement. Per-thread FP64 accumulation in keyline order, then the same shuffle tree / ticket scheme (1/KPT as many CTAs).
template <int KPT> __device__ __forceinline__ void tv_ld(const float* __restrict__ p, float* v) {
  if constexpr (KPT == 4) { const float4 a = *reinterpret_cast<const float4*>(p); v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; }
  else if constexpr (KPT == 2) { const float2 a = *reinterpret_cast<const float2*>(p); v[0] = a.x; v[1] = a.y; }
  else v[0] = *p;
}
template <int KPT> __device__ __forceinline__ void tv_ld(const unsigned* __restrict__ p, unsigned* v) {
  if constexpr (KPT == 4) { const uint4 a = *reinterpret_cast<const uint4*>(p); v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; }
  else if constexpr (KPT == 2) { const uint2 a = *reinterpret_cast<const uint2*>(p); v[0] = a.x; v[1] = a.y; }
  else v[0] = *p;
}
template <int KPT> __device__ __forceinline__ void tv_ld2(const float2* __restrict__ p, float2* v) {
  if constexpr (KPT >= 2) {
#pragma unroll
    for (int k = 0; k < KPT; k += 2) {
      const float4 a = *reinterpret_cast<const float4*>(p + k);
      v[k] = make_float2(a.x, a.y); v[k + 1] = make_float2(a.z, a.w);
    }
  } else v[0] = *p;
}
template <int KPT, typename T> __device__ __forceinline__ void tv_st(T* p, const T* v, int nvalid) {
  static_assert(sizeof(T) == 4, "32-bit elements");
  if (nvalid >= KPT) {
    const unsigned* u = reinterpret_cast<const unsigned*>(v);
    if constexpr (KPT == 4) *reinterpret_cast<uint4*>(p) = make_uint4(u[0], u[1], u[2], u[3]);
    else if constexpr (KPT == 2) *reinterpret_cast<uint2*>(p) = make_uint2(u[0], u[1]);
    else p[0] = v[0];
  } else {
#pragma unroll
    for (int k = 0; k < KPT; ++k) if (k < nvalid) p[k] = v[k];
  }
}

// ---- RBV_REDUCE_ORDERED addend layout (the chain kernels are further down, behind the FAST kernels) ----
#ifndef ORD_NST
#define ORD_NST 4             // stages per lane: 83 / 112 KB of shared memory, one chain CTA per SM anyway
#endif
#define ORD_PAD 4             // floats behind a row of a chunk: row r starts at bank 4 r, so the LDS.128 of different rows never collide
#ifndef TV_CH
#define TV_CH 512             // keylines per chunk: tryVel (10 rows) / extRotVel (27 rows). Per chunk the chain warp spends ~100 cycles on barriers
#endif
#ifndef ER_CH
#define ER_CH 256             // and addressing (measured 400 with 256 / 128-keyline chunks and the wait in line): small against CH x 4 cycles of adds
#endif
// Addend layout in HBM (c->prod): per lane, chunk after chunk; a chunk = NROWS rows of CH + ORD_PAD floats, i.e. exactly the image of one
// shared-memory stage, so ONE bulk copy (cp.async.bulk, 10.4 / 14.3 KB) per chunk stages it - issued by one thread of a loader warp, the chain
// warp never spends an issue slot on it. (Round 2 first kept [row][lane][keyline] planes and one copy per row and chunk from the chain warp
// itself: 10 - 27 serialised UBLKCP per chunk plus a proxy fence, 14 cycles per keyline instead of 4.)
template <int NROWS, int CH> struct OrdLayout {
  static constexpr int RS = CH + ORD_PAD, CHUNK = NROWS * RS;
  __host__ __device__ static size_t lane_stride(int kmax) { return (size_t)((kmax + CH - 1) / CH) * CHUNK; }
  __host__ __device__ static size_t at(int idx) { return (size_t)(idx / CH) * CHUNK + (size_t)(idx % CH); }   // + a * RS: accumulator a of keyline idx
};
typedef OrdLayout<10, TV_CH> OrdTV;
typedef OrdLayout<27, ER_CH> OrdER;
size_t tr_ordered_floats(int lanes, int kmax) {   // c->prod: the larger of the two layouts
  const size_t a = OrdTV::lane_stride(kmax), b = OrdER::lane_stride(kmax);
  return (size_t)lanes * (a > b ? a : b);
}
#define ORD_SMEM(L) (sizeof(float) * ORD_NST * L::CHUNK)

// ORD = RBV_REDUCE_ORDERED: instead of FP64 partial sums the block leaves the ten float32 addends of every keyline in ten planes
// (`rows`, zeros where the reference adds nothing: x + 0 == x exactly); k_chain_tryvel, launched behind it, adds them in keyline order.
template <int KPT, bool ORD>
__device__ __forceinline__ void tryvel_fastn_body(const int b, const int blk, KeylineSoA Ko, const int* __restrict__ count_o, const float* __restrict__ thr_o,
                                                         const float4* __restrict__ tgt, const unsigned* __restrict__ field, const float* __restrict__ vel_ptr,
                                                         int vel_stride, const float* __restrict__ sr_ptr, int sr_stride,
                                                         double* __restrict__ partials, unsigned* __restrict__ ticket,
                                                         const float* __restrict__ agg_prev, float* __restrict__ agg_cur, int nblk,
                                                         LMState* lm, float* lane_out, int lm_call, const DevParams& P,
                                                         float* __restrict__ rows, int* __restrict__ ord_flags = nullptr, bool persist = false) {
  __shared__ double sm[(TV_TPB / 32) * 10];
  __shared__ double result[10];
  __shared__ float warp_val[TV_TPB / 32];
  __shared__ int warp_has[TV_TPB / 32];
  __shared__ int s_last;
  constexpr int KPB = TV_TPB * KPT;
  const int n = count_o[b], t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int nact = max(1, (n + KPB - 1) / KPB);
  if (blk >= nact) return;
  const int idx0 = (blk * TV_TPB + t) * KPT;
  const size_t base = (size_t)b * P.kmax, o = base + idx0;
  const int nvalid = min(KPT, n - idx0);          // <= 0: nothing of this thread is a keyline
  double acc[10];
#pragma unroll
  for (int a = 0; a < 10; ++a) acc[a] = 0.0;
  float code[KPT], res[KPT];
#pragma unroll
  for (int k = 0; k < KPT; ++k) { code[k] = -2.0f; res[k] = 0.0f; }
  if (nvalid > 0) {
    // (__ldcg: in k_minimize_ord* the velocity was written by the lane's chain CTA earlier in the SAME launch - no stale L1 / read-only path)
    const float v0 = __ldcg(vel_ptr + (size_t)b * vel_stride), v1 = __ldcg(vel_ptr + (size_t)b * vel_stride + 1), v2 = __ldcg(vel_ptr + (size_t)b * vel_stride + 2);
    const float srmin = __ldcg(sr_ptr + (size_t)b * sr_stride), thr = thr_o[b];
    const unsigned* __restrict__ field_b = field + (size_t)b * P.npix;
    // ---- round trip 1: every field of the KPT keylines, no test in between
    float gn[KPT], srho[KPT], rho[KPT];
    unsigned mt[KPT];
    float2 pi[KPT], g[KPT];
    tv_ld<KPT>(Ko.gnorm + o, gn); tv_ld<KPT>(Ko.srho + o, srho); tv_ld<KPT>(Ko.rho + o, rho);
    if (lm_call != 0) {   // call 0 of minimizeVel starts from residuals = 0 (core.cpp:158): no fill pass, no read
      if (persist) {      // written by the previous call's block earlier in the same launch: read past L1
#pragma unroll
        for (int k = 0; k < KPT; ++k) res[k] = (k < nvalid) ? __ldcg(Ko.residual + o + k) : 0.0f;
      } else tv_ld<KPT>(Ko.residual + o, res);
    }
    tv_ld<KPT>(Ko.matches + o, mt); tv_ld2<KPT>(Ko.pos_img + o, pi); tv_ld2<KPT>(Ko.grad + o, g);
    float w[KPT], rho_p[KPT], px[KPT], py[KPT], pxc[KPT], pyc[KPT];
    bool skip[KPT], penalty[KPT];
    unsigned cell[KPT];
#pragma unroll
    for (int k = 0; k < KPT; ++k) {
      if (res[k] < 0.0f) res[k] = residual_lookback(agg_prev + (size_t)b * nblk, blk);
      skip[k] = (k >= nvalid) || (thr > 0.0f && gn[k] < thr) || (srho[k] > srmin) || (mt[k] < P.min_matches);   // core.cpp:88,91
      w[k] = 1.0f;
      if (res[k] > P.reweight_distance) w[k] = P.reweight_distance / res[k];           // :94
      const float z_p = (float)(1.0 / (double)rho[k] + (double)v2);                    // :96
      penalty[k] = z_p <= 0.0f;
      rho_p[k] = 0.0f; px[k] = py[k] = pxc[k] = pyc[k] = 0.0f;
      int x = 0, y = 0;
      if (!penalty[k]) {
        rho_p[k] = (float)(1.0 / (double)z_p);                                         // :104
        px[k] = rho_p[k] * (v0 * P.fm - v2 * pi[k].x) + pi[k].x;                       // :105
        py[k] = rho_p[k] * (v1 * P.fm - v2 * pi[k].y) + pi[k].y;
        pxc[k] = px[k] + P.cx; pyc[k] = py[k] + P.cy;
        x = rbv_d2i((double)pxc[k] + 0.5);                                             // :111
        y = rbv_d2i((double)pyc[k] + 0.5);
        penalty[k] = (x < 1 || y < 1 || x >= P.cols - 1 || y >= P.rows - 1);           // :114
      }
      // ---- round trip 2: the KPT field cells
      cell[k] = (!skip[k] && !penalty[k]) ? field_b[(size_t)y * P.cols + x] : RBV_FIELD_EMPTY;
    }
    // ---- round trip 3: the KPT target records (one 32-byte sector each)
    float4 t0[KPT], t1[KPT];
    int tid[KPT];
#pragma unroll
    for (int k = 0; k < KPT; ++k) {
      tid[k] = (int)(0xFFFFFFu - (cell[k] & 0xFFFFFFu));
      if (cell[k] != RBV_FIELD_EMPTY) { t0[k] = tgt[2 * (base + tid[k])]; t1[k] = tgt[2 * (base + tid[k]) + 1]; }
      else { t0[k] = make_float4(0.f, 0.f, 0.f, 0.f); t1[k] = t0[k]; }
    }
    int fwd[KPT];
    float oadd[ORD ? 10 : 1][KPT];   // ORD: the ten float32 addends of every keyline (zeros where the reference adds nothing: x + 0 == x)
#pragma unroll
    for (int k = 0; k < KPT; ++k) {
      fwd[k] = -1;
      if constexpr (ORD) {
#pragma unroll
        for (int a = 0; a < 10; ++a) oadd[a][k] = 0.0f;
      }
      if (skip[k]) continue;
      float add[10];
      if (penalty[k]) {
        const float f = (float)((1.0 / (double)srho[k]) * (double)P.search_range * (double)w[k]);   // :99 / :115
        if constexpr (ORD) oadd[0][k] = f * f;
        else acc[0] += (double)(f * f);
        continue;
      }
      float dfx = 0.0f, dfy = 0.0f, f;
      bool matched = false;
      if (cell[k] != RBV_FIELD_EMPTY) {
        const float n2 = gn[k] * gn[k];                                                // testfk :40-42
        const float dp = t0[k].x * g[k].x + t0[k].y * g[k].y;
        if (!(fabsf(dp - n2) > P.match_treshold * n2)) {
          matched = true;
          const float dx = pxc[k] - t0[k].z, dy = pyc[k] - t0[k].w;                    // :62-63
          const float gnx = t1[k].x, gny = t1[k].y;
          const float fi = dx * gnx + dy * gny;                                        // :67
          dfx = gnx / srho[k]; dfy = gny / srho[k];
          fwd[k] = tid[k];                                                             // :73
          f = fi / srho[k];                                                            // :75
          code[k] = fabsf(fi);                                                         // :141
        }
      }
      if (!matched) { f = P.search_range / srho[k]; code[k] = -1.0f; }                 // :52 / :59
      f *= w[k];
      const float jx = rho_p[k] * P.fm * dfx * w[k];                                   // :126-128
      const float jy = rho_p[k] * P.fm * dfy * w[k];
      const float jz = -rho_p[k] * (px[k] * dfx + py[k] * dfy) * w[k];
      add[0] = f * f;
      add[1] = jx * jx; add[2] = jy * jy; add[3] = jz * jz;
      add[4] = jx * jy; add[5] = jx * jz; add[6] = jy * jz;
      add[7] = jx * f; add[8] = jy * f; add[9] = jz * f;
#pragma unroll
      for (int a = 0; a < 10; ++a) {
        if constexpr (ORD) oadd[a][k] = add[a];
        else acc[a] += (double)add[a];
      }
    }
    tv_st<KPT, int>(Ko.match_fwd + o, fwd, nvalid);                                    // core.cpp:87
    if constexpr (ORD) {
      float* dst = rows + (size_t)b * OrdTV::lane_stride(P.kmax) + OrdTV::at(idx0);   // KPT divides TV_CH: the KPT keylines share a chunk row
#pragma unroll
      for (int a = 0; a < 10; ++a) tv_st<KPT, float>(dst + a * OrdTV::RS, oadd[a], nvalid);
    }
  }
  // carried |fi| in keyline order: thread aggregate = its last matched keyline, inclusive scan over the threads of the warp
  int has = 0; float val = 0.0f;
#pragma unroll
  for (int k = 0; k < KPT; ++k) if (code[k] >= 0.0f) { has = 1; val = code[k]; }
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const int h2 = __shfl_up_sync(0xFFFFFFFFu, has, off);
    const float v2 = __shfl_up_sync(0xFFFFFFFFu, val, off);
    if (lane >= off && !has) { has = h2; val = v2; }
  }
  if (lane == 31) { warp_has[warp] = has; warp_val[warp] = val; }
  int ehas = __shfl_up_sync(0xFFFFFFFFu, has, 1);                                      // exclusive: what enters this thread
  float eval = __shfl_up_sync(0xFFFFFFFFu, val, 1);
  if (lane == 0) { ehas = 0; eval = 0.0f; }
  if constexpr (!ORD) {
#pragma unroll
    for (int a = 0; a < 10; ++a) {
      double v = acc[a];
#pragma unroll
      for (int off = 16; off; off >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, off);
      if (lane == 0) sm[warp * 10 + a] = v;
    }
  }
  __syncthreads();
  if constexpr (ORD) {   // chain role of the same launch (k_tryvel_ord*, blockIdx.y == 0): its loader polls this flag before it stages the block's chunks.
    // The block's addends are ordered before the flag by the barrier above plus ONE cumulative gpu-scope fence + release store in thread 0
    // (the grid-sync pattern); a fence in every thread made all 256 of them wait for their stores to be acknowledged (`membar` stalls)
    // (blk * KPB < n: block 0 of a lane WITHOUT keylines also runs - it owns the lane's bookkeeping - but the chain has no chunk to
    // wait for then and would leave the flag standing for the next launch)
    if (ord_flags && !persist && t == 0 && blk * KPB < n) {
      __threadfence();
      asm volatile("st.release.gpu.global.s32 [%0], 1;" ::"l"(ord_flags + (size_t)b * nblk + blk) : "memory");
    }
  }
  if (!ehas || !has) {
    for (int w2 = warp - 1; w2 >= 0; --w2)
      if (warp_has[w2]) {
        if (!ehas) { ehas = 1; eval = warp_val[w2]; }
        if (!has) { has = 1; val = warp_val[w2]; }
        break;
      }
  }
  if (nvalid > 0) {
    float out[KPT];
#pragma unroll
    for (int k = 0; k < KPT; ++k) {
      if (code[k] >= 0.0f) { ehas = 1; eval = code[k]; }
      // core.cpp:141 (-1 = look back next time); no write this call = keep the (resolved) value that was read
      out[k] = code[k] >= -1.5f ? (ehas ? eval : (blk == 0 ? 0.0f : -1.0f)) : res[k];
    }
    tv_st<KPT, float>(Ko.residual + o, out, nvalid);
  }
  if (t == TV_TPB - 1) agg_cur[(size_t)b * nblk + blk] = has ? val : -1.0f;
  if constexpr (ORD) {
    if (persist) {   // k_minimize_ord*: the next call's block of these keylines reads the residuals and the block aggregate written above, so the
      // flag goes up behind them (barrier + one cumulative fence + release; the chain's release of the lane's call counter carries it on)
      __syncthreads();
      if (ord_flags && t == 0 && blk * KPB < n) {
        __threadfence();
        asm volatile("st.release.gpu.global.s32 [%0], 1;" ::"l"(ord_flags + (size_t)b * nblk + blk) : "memory");
      }
    }
    return;   // the addends are in `rows`: the chain adds them in keyline order
  }
  if (t < 10) {
    double s2 = 0.0;
#pragma unroll
    for (int w2 = 0; w2 < TV_TPB / 32; ++w2) s2 += sm[w2 * 10 + t];
    partials[((size_t)b * nblk + blk) * 10 + t] = s2;
  }
  __threadfence();
  __syncthreads();
  if (t == 0) s_last = (atomicAdd(&ticket[b], 1u) == (unsigned)(nact - 1));
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (t < 10) {
    double s2 = 0.0;
    for (int j = 0; j < nact; ++j) s2 += __ldcg(partials + ((size_t)b * nblk + j) * 10 + t);
    result[t] = s2;
  }
  __syncthreads();
  if (t == 0) {
    ticket[b] = 0u;
    float JtJ[9], JtF[3];
    const float score = (float)result[0];
    JtJ[0] = (float)result[1]; JtJ[4] = (float)result[2]; JtJ[8] = (float)result[3];
    JtJ[1] = JtJ[3] = (float)result[4]; JtJ[2] = JtJ[6] = (float)result[5]; JtJ[5] = JtJ[7] = (float)result[6];
    JtF[0] = (float)result[7]; JtF[1] = (float)result[8]; JtF[2] = (float)result[9];
    if (lm_call >= 0) {
      float vu[3] = {vel_ptr[(size_t)b * vel_stride], vel_ptr[(size_t)b * vel_stride + 1], vel_ptr[(size_t)b * vel_stride + 2]};
      lm_update(lm[b], lm_call, P.iterations, score, JtJ, JtF, vu);
    }
    float* outp = lane_out + (size_t)b * OUT_STRIDE;
    for (int i = 0; i < 9; ++i) outp[OUT_JTJ + i] = JtJ[i];
    for (int i = 0; i < 3; ++i) outp[OUT_JTF + i] = JtF[i];
    outp[OUT_SCORE] = score;
  }
}

#define RBV_TV_PARAMS KeylineSoA Ko, const int* __restrict__ count_o, const float* __restrict__ thr_o, const float4* __restrict__ tgt,             \
                      const unsigned* __restrict__ field, const float* __restrict__ vel_ptr, int vel_stride, const float* __restrict__ sr_ptr,    \
                      int sr_stride, double* __restrict__ partials, unsigned* __restrict__ ticket, const float* __restrict__ agg_prev,            \
                      float* __restrict__ agg_cur, int nblk, LMState* lm, float* lane_out, int lm_call, const __grid_constant__ DevParams P
#define RBV_TV_PARAMS_DEV KeylineSoA Ko, const int* __restrict__ count_o, const float* __restrict__ thr_o, const float4* __restrict__ tgt,         \
                          const unsigned* __restrict__ field, const float* __restrict__ vel_ptr, int vel_stride, const float* __restrict__ sr_ptr, \
                          int sr_stride, double* __restrict__ partials, unsigned* __restrict__ ticket, const float* __restrict__ agg_prev,         \
                          float* __restrict__ agg_cur, int nblk, LMState* lm, float* lane_out, int lm_call, const DevParams& P
#define RBV_TV_PASS Ko, count_o, thr_o, tgt, field, vel_ptr, vel_stride, sr_ptr, sr_stride, partials, ticket, agg_prev, agg_cur, nblk, lm, lane_out, lm_call, P
__global__ void __launch_bounds__(TV_TPB) k_tryvel_fast2(RBV_TV_PARAMS) { tryvel_fastn_body<2, false>(blockIdx.y, blockIdx.x, RBV_TV_PASS, nullptr); }
__global__ void __launch_bounds__(TV_TPB, TV_MINB) k_tryvel_fast4(RBV_TV_PARAMS) { tryvel_fastn_body<4, false>(blockIdx.y, blockIdx.x, RBV_TV_PASS, nullptr); }

// ---- RBV_REDUCE_ORDERED: the reference's float32 sums in keyline order --------------------------------------------------------
// The reference adds its ten tryVel accumulators (and the 27 of extRotVel) sequentially in float32 over the keylines
// (core.cpp:130-139,244-245); match indices downstream depend on those bits. A sequential float32 sum cannot be re-associated,
// so the chain itself stays serial: ONE warp per lane, thread a = accumulator a, one dependent FADD (4 cycles, ~2 ns) per keyline.
// What can be done is to take everything else off that chain:
//   * the evaluation blocks (k_tryvel_ord2/4, k_extrot_ord: the FAST kernels' per-keyline arithmetic) leave the float32 ADDENDS of every
//     keyline in the chunked layout above (OrdLayout; zeros where the reference adds nothing: x + 0 == x exactly) and raise a flag per block;
//   * one chain CTA per lane rides in the SAME launch (grid row 0, dispatched first; 64 lanes = 64 SMs, the rest of the GPU stays free):
//     its loader warp takes the flags of the blocks a chunk covers and stages the chunk with ONE TMA bulk copy into a 4-deep shared-memory
//     ring (mbarrier completion), its chain warp adds row a in lane a: nothing but an LDS.128 per four keylines and the FADDs on that
//     warp. (Forming the products on the chain warp costs an FMUL per keyline next to the FADD; the FP32 pipe takes a 3-register
//     instruction every 2 cycles per scheduler, and the chain ran at 8 cycles per keyline instead of 4.)
// RBV_TUNE_ORDERED = 2 runs the same chain as a kernel of its own behind the evaluation (k_chain_tryvel / k_chain_extrot); = 4 keeps the
// chain CTAs for all tryVel calls of a minimisation (k_minimize_ord*). History of the round (wrong sums that turned out to be an nvcc
// unroller bug in the block loop, the staging forms tried): DESIGN.md "ORDERED reductions".

// sixteen float32 adds in keyline order
__device__ __forceinline__ void ord_add16(float& s, const float4& a, const float4& b, const float4& c, const float4& d) {
  s += a.x; s += a.y; s += a.z; s += a.w; s += b.x; s += b.y; s += b.z; s += b.w;
  s += c.x; s += c.y; s += c.z; s += c.w; s += d.x; s += d.y; s += d.z; s += d.w;
}
// CTA = 2 warps. Warp 1 (one thread) is the loader: waits for a free stage, arms its `full` barrier with the chunk's bytes and issues the
// bulk copy. Warp 0 is the chain: lane a adds row a of every staged chunk in keyline order, then releases the stage (`empty`).
// Returns the sum in lanes < NROWS of warp 0; warp 1 returns 0 and has nothing left to do.
// `flags` (chain role inside the evaluation launch): flags[j] goes to 1 when evaluation block j (kpb keylines) has left its addends; the
// loader takes (and clears) the flags of the blocks a chunk covers before it copies the chunk. nullptr: the addends are complete (kernel boundary).
template <typename L>
__device__ __forceinline__ float ordered_chain(const float* __restrict__ lane_rows, int n, float* ring,
                                               unsigned long long* bars /* [2 * ORD_NST], shared: full, empty */, float* dbg = nullptr,
                                               int* flags = nullptr, int kpb = 1) {
  constexpr int NROWS = L::CHUNK / L::RS, CH = L::RS - ORD_PAD, RS = L::RS, CHUNK = L::CHUNK;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nch = (n + CH - 1) / CH;
  unsigned long long* full = bars;
  unsigned long long* empty = bars + ORD_NST;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 2 * ORD_NST; ++i) mbar_init(bars + i, 1u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("bar.sync 1, 64;" ::: "memory");   // the two warps of the chain (the CTA may have more threads: they have left)
  float s = 0.0f;
  if (warp == 1) {
    if (lane == 0) {
      int have = 0;
#pragma unroll 1
      for (int c = 0; c < nch; ++c) {
        const int stage = c % ORD_NST;
        if (flags) {
          const int need = (min(n, (c + 1) * CH) - 1) / kpb + 1;   // evaluation blocks up to the last keyline of this chunk
          for (; have < need; ++have) {
            int v;
            long long t0 = 0;
            for (int spin = 0;; ++spin) {
              asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flags + have) : "memory");
              if (v) break;
              if ((spin & 1023) == 1023) {
                if (t0 == 0) t0 = clock64();
                else if (clock64() - t0 > 8000000000ll) __trap();   // seconds: a protocol bug must fault, never hang the device
              }
            }
            flags[have] = 0;   // for the next launch (stream order)
          }
          asm volatile("fence.proxy.async;" ::: "memory");   // the blocks' generic-proxy stores before this thread's async-proxy (bulk copy) reads
        }
        if (c >= ORD_NST) mbar_wait(empty + stage, (unsigned)((c / ORD_NST - 1) & 1));   // the chain has released this stage's previous chunk
        mbar_arrive_expect_tx(full + stage, (unsigned)(CHUNK * 4));
        bulk_g2s(ring + (size_t)stage * CHUNK, lane_rows + (size_t)c * CHUNK, (unsigned)(CHUNK * 4), full + stage);
      }
    }
    return 0.0f;
  }
  bool ready = false;
#pragma unroll 1
  for (int c = 0; c < nch; ++c) {
    const int stage = c % ORD_NST;
    if (!ready) mbar_wait(full + stage, (unsigned)((c / ORD_NST) & 1));
    // the next chunk's barrier is probed now and the answer used after this chunk's adds: the loader runs stages ahead, so the probe
    // nearly always says yes and its ~150-cycle latency never sits in the chain (a no falls back to the blocking wait above)
    ready = (c + 1 < nch) && mbar_test(full + (c + 1) % ORD_NST, (unsigned)(((c + 1) / ORD_NST) & 1));
    const int cnt = min(CH, n - c * CH);
    if (lane < NROWS) {   // lane a adds accumulator a: nothing but LDS.128 and the dependent FADDs on this warp (the FP32 pipe takes a 3-register
      // instruction every 2 cycles per scheduler: one FMUL per keyline next to the FADD made the 4-cycle chain run at 8)
      const float4* ra = reinterpret_cast<const float4*>(ring + (size_t)stage * CHUNK + lane * RS);
      const int n16 = cnt >> 4;
      // two 16-keyline blocks per trip, register sets A and B, next block loaded while the current one is added.
      // unroll 1 on purpose: nvcc 12.9 unrolled the one-block form of this loop by 8 with 4 / 2 / 1-block remainder sections, and in those the
      // prefetch for the NEXT section was addressed from the already advanced block counter plus the old offset (block b + 8 instead of b + 4):
      // chunks with (n16 & 7) in {3, 5, 6, 7} added a block from beyond the chunk (found with RBV_DEBUG_CHAIN, tools/debug_ordered.py).
      float4 A0, A1, A2, A3, B0, B1, B2, B3;
      if (n16 > 0) { A0 = ra[0]; A1 = ra[1]; A2 = ra[2]; A3 = ra[3]; }
      int b16 = 0;
#pragma unroll 1
      for (; b16 + 2 <= n16; b16 += 2) {
        const float4* rb = ra + 4 * b16;
        B0 = rb[4]; B1 = rb[5]; B2 = rb[6]; B3 = rb[7];
        ord_add16(s, A0, A1, A2, A3);
        if (b16 + 2 < n16) { A0 = rb[8]; A1 = rb[9]; A2 = rb[10]; A3 = rb[11]; }
        ord_add16(s, B0, B1, B2, B3);
      }
      if (b16 < n16) ord_add16(s, A0, A1, A2, A3);
      const float* rt = reinterpret_cast<const float*>(ra);
#pragma unroll 1
      for (int i = n16 * 16; i < cnt; ++i) s += rt[i];
    }
    if (dbg && lane == 0) dbg[c] = s;   // debug (RBV_DEBUG_CHAIN): running sum of accumulator 0 after every chunk
    __syncwarp();
    if (lane == 0) mbar_arrive(empty + stage);
  }
  return s;
}

// warp 0 of a chain: the ten sums -> LM step (core.cpp:159-185) and the lane's outputs
__device__ __forceinline__ void tryvel_chain_finish(float s, int b, LMState* lm, float* lane_out, const float* __restrict__ vel_ptr, int vel_stride,
                                                    int lm_call, const DevParams& P) {
  float r[10];
#pragma unroll
  for (int a = 0; a < 10; ++a) r[a] = __shfl_sync(0xFFFFFFFFu, s, a);
  if ((threadIdx.x & 31) == 0) {
    float JtJ[9], JtF[3];
    const float score = r[0];
    JtJ[0] = r[1]; JtJ[4] = r[2]; JtJ[8] = r[3];
    JtJ[1] = JtJ[3] = r[4]; JtJ[2] = JtJ[6] = r[5]; JtJ[5] = JtJ[7] = r[6];
    JtF[0] = r[7]; JtF[1] = r[8]; JtF[2] = r[9];
    if (lm_call >= 0) {
      float vu[3] = {vel_ptr[(size_t)b * vel_stride], vel_ptr[(size_t)b * vel_stride + 1], vel_ptr[(size_t)b * vel_stride + 2]};
      lm_update(lm[b], lm_call, P.iterations, score, JtJ, JtF, vu);
    }
    float* outp = lane_out + (size_t)b * OUT_STRIDE;
    for (int i = 0; i < 9; ++i) outp[OUT_JTJ + i] = JtJ[i];
    for (int i = 0; i < 3; ++i) outp[OUT_JTF + i] = JtF[i];
    outp[OUT_SCORE] = score;
  }
}
// Evaluation + chain in ONE launch (ord_flags != nullptr): grid (lanes, 1 + blocks). Row y = 0 of the grid is dispatched first: one chain CTA
// per lane (2 of its warps stay: chain + loader); rows y >= 1 are the evaluation blocks, block index fastest over the lanes, so every lane's
// chain finds its first chunks after the first wave and the adds run behind the evaluation instead of after it. The evaluation never waits
// for a chain, so any dispatch order terminates. ord_flags == nullptr: evaluation only, grid (lanes, blocks), k_chain_tryvel behind it.
template <int KPT>
__device__ __forceinline__ void tryvel_ord_entry(RBV_TV_PARAMS_DEV, float* __restrict__ rows, int* __restrict__ ord_flags, float* dbg) {
  extern __shared__ __align__(128) float ord_ring[];
  __shared__ __align__(8) unsigned long long bars[2 * ORD_NST];
  const int b = blockIdx.x;
  if (ord_flags && blockIdx.y == 0) {
    if (threadIdx.x >= 64) return;
    const float s = ordered_chain<OrdTV>(rows + (size_t)b * OrdTV::lane_stride(P.kmax), count_o[b], ord_ring, bars,
                                         dbg ? dbg + (size_t)b * P.kmax : nullptr, ord_flags + (size_t)b * nblk, TV_TPB * KPT);
    if (threadIdx.x < 32) tryvel_chain_finish(s, b, lm, lane_out, vel_ptr, vel_stride, lm_call, P);
    return;
  }
  tryvel_fastn_body<KPT, true>(b, (int)blockIdx.y - (ord_flags ? 1 : 0), RBV_TV_PASS, rows, ord_flags);
}
__global__ void __launch_bounds__(TV_TPB, TV_MINB) k_tryvel_ord2(RBV_TV_PARAMS, float* __restrict__ rows, int* __restrict__ ord_flags, float* dbg) {
  tryvel_ord_entry<2>(RBV_TV_PASS, rows, ord_flags, dbg);
}
__global__ void __launch_bounds__(TV_TPB, TV_MINB) k_tryvel_ord4(RBV_TV_PARAMS, float* __restrict__ rows, int* __restrict__ ord_flags, float* dbg) {
  tryvel_ord_entry<4>(RBV_TV_PASS, rows, ord_flags, dbg);
}
// Core::minimizeVel in ONE launch (RBV_TUNE_ORDERED = 4): grid (lanes, 1 + calls x blocks). Row 0: one chain CTA per lane, which stays for all
// iterations + 1 tryVel calls: chain, LM step, then it publishes the lane's call counter (lm_done, reset by k_quantile in front of every
// minimisation). Rows 1..: the evaluation blocks of call 0, then of call 1, ...; a block of call t waits for lm_done[lane] >= t, i.e. for the
// velocity its call evaluates. Lanes never wait for each other: a lane with few keylines runs ahead, and no call pays for the slowest
// lane's chain as the per-call launches do. Progress: blocks are dispatched in linear order, so every block another block can wait for
// (the lane's chain CTA, the blocks of earlier calls) is resident or finished by then; a waiting block only holds its slot.
template <int KPT>
__device__ __forceinline__ void minimize_ord_entry(KeylineSoA Ko, const int* __restrict__ count_o, const float* __restrict__ thr_o,
                                                   const float4* __restrict__ tgt, const unsigned* __restrict__ field, float* blkagg, int parity0,
                                                   int nblk, LMState* lm, float* lane_out, int ncall, const DevParams& P, float* __restrict__ rows,
                                                   int* __restrict__ ord_flags, int* lm_done, int nb) {
  extern __shared__ __align__(128) float ord_ring[];
  __shared__ __align__(8) unsigned long long bars[2 * ORD_NST];
  const int b = blockIdx.x;
  const float* lmf = reinterpret_cast<const float*>(lm);
  constexpr int LMF = (int)(sizeof(LMState) / sizeof(float));
  const float* vel0 = lmf + offsetof(LMState, vel) / sizeof(float);
  const float* vel1 = lmf + offsetof(LMState, vnew) / sizeof(float);
  if (blockIdx.y == 0) {
    if (threadIdx.x >= 64) return;
    for (int t = 0; t < ncall; ++t) {
      const float s = ordered_chain<OrdTV>(rows + (size_t)b * OrdTV::lane_stride(P.kmax), count_o[b], ord_ring, bars, nullptr,
                                           ord_flags + (size_t)b * nblk, TV_TPB * KPT);
      if (threadIdx.x < 32) {
        tryvel_chain_finish(s, b, lm, lane_out, t == 0 ? vel0 : vel1, LMF, t, P);
        if (threadIdx.x == 0) {   // the LM state of call t is out: the lane's blocks of call t + 1 may go
          __threadfence();
          asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(lm_done + b), "r"(t + 1) : "memory");
        }
      }
    }
    return;
  }
  const int t = ((int)blockIdx.y - 1) / nb, blk = ((int)blockIdx.y - 1) % nb;
  {
    const int n = count_o[b];
    if (blk >= max(1, (n + TV_TPB * KPT - 1) / (TV_TPB * KPT))) return;   // before any waiting: nothing of this block is a keyline
  }
  if (t > 0) {
    if (threadIdx.x == 0) {
      long long t0 = 0;
      for (int spin = 0;; ++spin) {
        int v;
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(lm_done + b) : "memory");
        if (v >= t) break;
        __nanosleep(200);
        if ((spin & 1023) == 1023) {
          if (t0 == 0) t0 = clock64();
          else if (clock64() - t0 > 8000000000ll) __trap();   // seconds: a protocol bug must fault, never hang the device
        }
      }
    }
    __syncthreads();
  }
  const size_t half = (size_t)P.lanes * nblk;
  const int par = (parity0 + t) & 1;
  tryvel_fastn_body<KPT, true>(b, blk, Ko, count_o, thr_o, tgt, field, t == 0 ? vel0 : vel1, LMF, lmf + offsetof(LMState, sigma_rho_min) / sizeof(float), LMF,
                               nullptr, nullptr, blkagg + (size_t)(1 - par) * half, blkagg + (size_t)par * half, nblk, lm, lane_out, t, P, rows, ord_flags, true);
}
#define RBV_MIN_PARAMS KeylineSoA Ko, const int* __restrict__ count_o, const float* __restrict__ thr_o, const float4* __restrict__ tgt,          \
                       const unsigned* __restrict__ field, float* blkagg, int parity0, int nblk, LMState* lm, float* lane_out, int ncall,        \
                       const __grid_constant__ DevParams P, float* __restrict__ rows, int* __restrict__ ord_flags, int* lm_done, int nb
#define RBV_MIN_PASS Ko, count_o, thr_o, tgt, field, blkagg, parity0, nblk, lm, lane_out, ncall, P, rows, ord_flags, lm_done, nb
__global__ void __launch_bounds__(TV_TPB) k_minimize_ord2(RBV_MIN_PARAMS) { minimize_ord_entry<2>(RBV_MIN_PASS); }
__global__ void __launch_bounds__(TV_TPB) k_minimize_ord4(RBV_MIN_PARAMS) { minimize_ord_entry<4>(RBV_MIN_PASS); }

// the two-kernel form (RBV_TUNE_ORDERED = 2): the ten sums of one tryVel call in keyline order + the LM step, behind the evaluation kernel
__global__ void __launch_bounds__(64) k_chain_tryvel(const int* __restrict__ count_o, const float* __restrict__ rows, LMState* lm, float* lane_out,
                                                     const float* __restrict__ vel_ptr, int vel_stride, int lm_call,
                                                     const __grid_constant__ DevParams P, float* dbg) {
  extern __shared__ __align__(128) float ord_ring[];
  __shared__ __align__(8) unsigned long long bars[2 * ORD_NST];
  const int b = blockIdx.x;
  const float s = ordered_chain<OrdTV>(rows + (size_t)b * OrdTV::lane_stride(P.kmax), count_o[b], ord_ring, bars, dbg ? dbg + (size_t)b * P.kmax : nullptr);
  if (threadIdx.x < 32) tryvel_chain_finish(s, b, lm, lane_out, vel_ptr, vel_stride, lm_call, P);
}

__global__ void __launch_bounds__(TV_TPB) k_residual_resolve(KeylineSoA Ko, const int* __restrict__ count_o, const float* __restrict__ agg,
                                                             int nblk, int kmax, int kpb) {
  const int b = blockIdx.y, idx = blockIdx.x * TV_TPB + threadIdx.x;
  if (idx >= count_o[b]) return;
  const size_t o = (size_t)b * kmax + idx;
  if (Ko.residual[o] < 0.0f) Ko.residual[o] = residual_lookback(agg + (size_t)b * nblk, idx / kpb);
}

// RBV_REDUCE_ORDERED through the first implementation (k_tryvel_eval + k_tryvel_reduce, k_extrot_rows + k_extrot_reduce: evaluation, then
// one block per lane adding behind it) instead of the chain warps next to the evaluation; kept as a cross-check (RBV_TUNE_ORDERED = 1)
int tr_ordered_legacy(const rbv_ctx* c) { return c->reduce_mode == RBV_REDUCE_ORDERED && c->tune[RBV_TUNE_ORDERED] == 1; }

// all iterations + 1 tryVel calls of Core::minimizeVel in one launch (k_minimize_ord*); 0 = not this configuration, the caller launches call by call
int tr_minimize_persist_wanted(const rbv_ctx* c) {
  return c->reduce_mode == RBV_REDUCE_ORDERED && c->tune[RBV_TUNE_ORDERED] == 4 && c->prof_mode == -1 ? 1 : 0;
}
int tr_minimize_persist(rbv_ctx* c, rbv_map* map) {
  const DevParams& P = c->P;
  if (!c->field_map) { rbv_set_error("minimizeVel before buildDistanceField"); return RBV_ERR_STATE; }
  int kpt = c->tune[RBV_TUNE_TRYVEL] ? c->tune[RBV_TUNE_TRYVEL] : (P.lanes >= 8 ? 4 : 2);
  if (kpt != 2) kpt = 4;
  const int nb = (P.kmax + TV_TPB * kpt - 1) / (TV_TPB * kpt), ncall = P.iterations + 1;
  const dim3 grid(P.lanes, 1 + ncall * nb);
  if (grid.y > 65535u) { rbv_set_error("minimizeVel: too many blocks for one launch"); return RBV_ERR_ARG; }
  if (kpt == 2) { RBV_LAUNCH(c, k_minimize_ord2, (grid), (TV_TPB), ORD_SMEM(OrdTV), map->k, map->count, map->threshold, c->tgt, c->field, c->blkagg, c->tv_parity, c->nblk, c->lm, c->lane_out, ncall, P, c->prod, c->ord_flags, c->lm_done, nb); }
  else { RBV_LAUNCH(c, k_minimize_ord4, (grid), (TV_TPB), ORD_SMEM(OrdTV), map->k, map->count, map->threshold, c->tgt, c->field, c->blkagg, c->tv_parity, c->nblk, c->lm, c->lane_out, ncall, P, c->prod, c->ord_flags, c->lm_done, nb); }
  const size_t half = (size_t)P.lanes * c->nblk;
  float* agg_last = c->blkagg + (size_t)((c->tv_parity + ncall - 1) & 1) * half;   // the aggregates of the last call (residual look-back of pending keylines)
  if (c->seq_fused) { c->seq_resolve_agg = agg_last; c->seq_resolve_kpb = TV_TPB * kpt; }
  else RBV_LAUNCH(c, k_residual_resolve, (dim3(c->nblk, P.lanes)), (TV_TPB), 0, map->k, map->count, agg_last, c->nblk, P.kmax, TV_TPB * kpt);
  c->tv_parity = (c->tv_parity + ncall) & 1;
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// lm_call = -1: single evaluation at lane_in vel / sigma_rho_min; 0: first LM call at lm.vel; >= 1: at lm.vnew
int tr_try_vel(rbv_ctx* c, rbv_map* map, int lm_call) {
  const DevParams& P = c->P;
  if (!c->field_map) { rbv_set_error("tryVel before buildDistanceField"); return RBV_ERR_STATE; }
  const float* vel_ptr; int vel_stride; const float* sr_ptr; int sr_stride;
  const int lm_floats = (int)(sizeof(LMState) / sizeof(float));
  if (lm_call < 0) {
    vel_ptr = c->lane_in + IN_VEL; vel_stride = IN_STRIDE; sr_ptr = c->lane_in + IN_SRMIN; sr_stride = IN_STRIDE;
  } else {
    const float* lmf = reinterpret_cast<const float*>(c->lm);
    vel_ptr = lmf + (lm_call == 0 ? offsetof(LMState, vel) : offsetof(LMState, vnew)) / sizeof(float);
    vel_stride = lm_floats;
    sr_ptr = lmf + offsetof(LMState, sigma_rho_min) / sizeof(float);
    sr_stride = lm_floats;
  }
  if (tr_ordered_legacy(c)) {
    RBV_LAUNCH(c, k_tryvel_eval, (dim3((P.keylines_max + 255) / 256, P.lanes)), (256), 0, map->k, map->count, map->threshold, c->tgt, c->field, vel_ptr, vel_stride, sr_ptr, sr_stride, c->prod, c->fiabs, P);
    RBV_LAUNCH(c, k_tryvel_reduce, (P.lanes), (1024), 0, map->k, map->count, c->prod, c->fiabs, c->lm, c->lane_out, vel_ptr, vel_stride, lm_call, 1, P);
  } else {
    const bool ord = c->reduce_mode == RBV_REDUCE_ORDERED;
    const int nblk = c->nblk;
    const size_t half = (size_t)P.lanes * nblk;
    float* agg_cur = c->blkagg + (size_t)c->tv_parity * half;
    const float* agg_prev = c->blkagg + (size_t)(1 - c->tv_parity) * half;
    // keylines per thread (RBV_TUNE_TRYVEL): 1 = k_tryvel_fast (FAST only), 2 / 4 = k_tryvel_{fast,ord}2 / 4 (0 = automatic: 4, or 2 with fewer than 8 lanes)
    int kpt = c->tune[RBV_TUNE_TRYVEL] ? c->tune[RBV_TUNE_TRYVEL] : (P.lanes >= 8 ? 4 : 2);   // few lanes: more, smaller CTAs
    if (kpt == 3 || kpt > 4) kpt = 4;
    if (ord && kpt == 1) kpt = 2;
    const int nb = (P.kmax + TV_TPB * kpt - 1) / (TV_TPB * kpt);
#define RBV_TV_ARGS map->k, map->count, map->threshold, c->tgt, c->field, vel_ptr, vel_stride, sr_ptr, sr_stride, c->partials, c->ticket, agg_prev, agg_cur, nblk, c->lm, c->lane_out, lm_call, P
    if (ord) {   // evaluation (addends to c->prod) with the chain CTAs in the same launch, or (RBV_TUNE_ORDERED = 2) the chain kernel behind it
      const bool two = c->tune[RBV_TUNE_ORDERED] == 2;
      float* dbg = getenv("RBV_DEBUG_CHAIN") ? c->fiabs : nullptr;
      int* flags = two ? nullptr : c->ord_flags;
      const dim3 grid(P.lanes, nb + (two ? 0 : 1));
      const size_t smem = two ? 0 : ORD_SMEM(OrdTV);
      if (kpt == 2) { RBV_LAUNCH(c, k_tryvel_ord2, (grid), (TV_TPB), smem, RBV_TV_ARGS, c->prod, flags, dbg); }
      else { RBV_LAUNCH(c, k_tryvel_ord4, (grid), (TV_TPB), smem, RBV_TV_ARGS, c->prod, flags, dbg); }
      if (two) RBV_LAUNCH(c, k_chain_tryvel, (P.lanes), (64), ORD_SMEM(OrdTV), map->count, c->prod, c->lm, c->lane_out, vel_ptr, vel_stride, lm_call, P, dbg);
    }
    else if (kpt == 1) { RBV_LAUNCH(c, k_tryvel_fast, (dim3(nb, P.lanes)), (TV_TPB), 0, RBV_TV_ARGS); }
    else if (kpt == 2) { RBV_LAUNCH(c, k_tryvel_fast2, (dim3(nb, P.lanes)), (TV_TPB), 0, RBV_TV_ARGS); }
    else { RBV_LAUNCH(c, k_tryvel_fast4, (dim3(nb, P.lanes)), (TV_TPB), 0, RBV_TV_ARGS); }
#undef RBV_TV_ARGS
    if (c->seq_fused && lm_call == P.iterations) {   // rbv_track_g1: k_fwd_link (next over the same keylines) clears the markers
      c->seq_resolve_agg = agg_cur; c->seq_resolve_kpb = TV_TPB * kpt;
    } else if (lm_call < 0 || lm_call == P.iterations)
      RBV_LAUNCH(c, k_residual_resolve, (dim3(nblk, P.lanes)), (TV_TPB), 0, map->k, map->count, agg_cur, nblk, P.kmax, TV_TPB * kpt);
    c->tv_parity = 1 - c->tv_parity;
  }
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// forwardMatch (edge_map.cpp:73-99). The sequential loop visits old keylines in index order and lets a later one
// overwrite the target unless the target already holds a larger rho; the count includes overwrites. Parallel form:
// (a) every old keyline with a forward match pushes itself on a per-target list (atomicExch), (b) one thread per
// target replays its (short) list in ascending index order with the reference's rule, so result and count are exact.
__global__ void k_fwd_link(KeylineSoA Ko, const int* __restrict__ count_o, int* __restrict__ head, int* __restrict__ next, int kmax,
                           const float* __restrict__ agg, int nblk, int kpb) {
  const int b = blockIdx.y, idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count_o[b]) return;
  const size_t base = (size_t)b * kmax;
  if (agg) {   // rbv_track_g1: k_residual_resolve's job rides along (same thread-per-old-keyline shape, one launch less)
    const size_t o = base + idx;
    if (Ko.residual[o] < 0.0f) Ko.residual[o] = residual_lookback(agg + (size_t)b * nblk, idx / kpb);
  }
  const int f = Ko.match_fwd[base + idx];
  if (f < 0) return;
  next[base + idx] = atomicExch(&head[base + f], idx);
}

__global__ void k_fwd_resolve(KeylineSoA Ko, KeylineSoA Kn, const int* __restrict__ count_n, int* __restrict__ head,
                              const int* __restrict__ next, int* __restrict__ lane_cnt, int kmax) {
  const int b = blockIdx.y, f = blockIdx.x * blockDim.x + threadIdx.x;
  int writes = 0;
  if (f < count_n[b]) {
    const size_t base = (size_t)b * kmax;
    const int h = head[base + f];
    if (h >= 0) {
      head[base + f] = -1;   // consumed: the head table is left empty for the next frame (no fill launch per frame)
      int match_id = Kn.match_id[base + f];
      float rho = Kn.rho[base + f];
      int last = -1, winner = -1;
      for (;;) {
        int cur = 0x7FFFFFFF;                       // smallest list entry > last
        for (int e = h; e >= 0; e = next[base + e])
          if (e > last && e < cur) cur = e;
        if (cur == 0x7FFFFFFF) break;
        last = cur;
        const float krho = Ko.rho[base + cur];
        if (match_id >= 0 && rho > krho) continue;                                       // :83
        rho = krho; match_id = cur; winner = cur;
        ++writes;
      }
      if (winner >= 0) {                                                                 // :86-93
        const size_t w = base + winner, o = base + f;
        Kn.rho[o] = rho;
        Kn.srho[o] = Ko.srho[w];
        Kn.matches[o] = Ko.matches[w] + 1u;
        Kn.match_id[o] = winner;
        Kn.match_pos_img[o] = Ko.pos_img[w];
        Kn.match_grad[o] = Ko.grad[w];
        Kn.match_gnorm[o] = Ko.gnorm[w];
        Kn.match_kf[o] = Ko.match_kf[w];
      }
    }
  }
  // count: warp-aggregated integer atomics (order independent)
#pragma unroll
  for (int off = 16; off; off >>= 1) writes += __shfl_down_sync(0xFFFFFFFFu, writes, off);
  if ((threadIdx.x & 31) == 0 && writes) atomicAdd(&lane_cnt[b * CNT_STRIDE + CNT_FWD], writes);
}

__global__ void k_zero_cnt(int* lane_cnt, int slot, unsigned long long* probes) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  lane_cnt[b * CNT_STRIDE + slot] = 0;
  if (slot == CNT_DIR) { lane_cnt[b * CNT_STRIDE + CNT_KF] = 0; if (probes) probes[b] = 0ull; }
}

int tr_forward_match(rbv_ctx* c, rbv_map* old_map, rbv_map* new_map) {
  const DevParams& P = c->P;
  // c->fwd_head is all -1 here: filled at creation, and k_fwd_resolve empties every list it replays
  if (!c->seq_fused) RBV_LAUNCH(c, k_zero_cnt, (1), (P.lanes), 0, c->lane_cnt, CNT_FWD, nullptr);
  dim3 g((P.keylines_max + 255) / 256, P.lanes);
  RBV_LAUNCH(c, k_fwd_link, (g), (256), 0, old_map->k, old_map->count, c->fwd_head, c->fwd_next, P.kmax,
             c->seq_resolve_agg, c->nblk, c->seq_resolve_kpb);
  c->seq_resolve_agg = nullptr;
  RBV_LAUNCH(c, k_fwd_resolve, (g), (256), 0, old_map->k, new_map->k, new_map->count, c->fwd_head, c->fwd_next, c->lane_cnt, P.kmax);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// extRotVel rows (core.cpp:202-236): Phi_j (6) and Y_j after the division by s_y*weight; zeros for unmatched keylines.
// what a row reads of its keyline; fetched with no test in between so that all loads of a thread are in flight together
struct ExtRotIn {
  int mid;
  float2 g, m, q;
  float gn, rho, sr;
};
__device__ __forceinline__ ExtRotIn extrot_load(const KeylineSoA& K, size_t o) {
  ExtRotIn in;
  in.mid = K.match_id[o]; in.g = K.grad[o]; in.gn = K.gnorm[o]; in.rho = K.rho[o];
  in.m = K.match_pos_img[o]; in.q = K.pos_img[o]; in.sr = K.srho[o];
  return in;
}
__device__ __forceinline__ void extrot_row_in(const ExtRotIn& in, float v0, float v1, float v2, const DevParams& P, float* phi) {
#pragma unroll
  for (int a = 0; a < 7; ++a) phi[a] = 0.0f;
  if (in.mid < 0) return;
  const float fm = P.fm;
  const float2 g = in.g;
  const float gn = in.gn;
  const float ux = g.x / gn, uy = g.y / gn;                                            // :207-208
  const float rho_t = (float)(1.0 / (1.0 / (double)in.rho + (double)v2));              // :210
  const float2 m = in.m;
  const float qtx = m.x + rho_t * (v0 * fm - v2 * m.x);                                // :211-212
  const float qty = m.y + rho_t * (v1 * fm - v2 * m.y);
  const float2 q = in.q;
  phi[0] = ux * rho_t * fm;                                                            // :217-222
  phi[1] = uy * rho_t * fm;
  phi[2] = ux * (-rho_t * q.x) + uy * (-rho_t * q.y);
  phi[3] = -ux * q.x * q.y / fm - uy * (fm + q.y * q.y / fm);
  phi[4] = uy * q.x * q.y / fm + ux * (fm + q.x * q.x / fm);
  phi[5] = -ux * q.y + uy * q.x;
  float y = ux * (q.x - qtx) + uy * (q.y - qty);                                       // :224
  const float dqvel = ux * (v0 * fm - v2 * m.x) + uy * (v1 * fm - v2 * m.y);           // :226
  const float sr = in.sr;
  const float sy = sqrtf(sr * sr * dqvel * dqvel + P.pixel_uncertainty * P.pixel_uncertainty);   // :227
  float w = 1.0f;
  if (fabsf(y) > P.reweight_distance) w = fabsf(y) / P.reweight_distance;              // :230
  const float dv = sy * w;
#pragma unroll
  for (int a = 0; a < 6; ++a) phi[a] /= dv;                                            // :232
  y /= dv;                                                                             // :233
  phi[6] = y;
}
__device__ __forceinline__ void extrot_row(const KeylineSoA& K, size_t o, float v0, float v1, float v2, const DevParams& P, float* phi) {
  if (K.match_id[o] < 0) {
#pragma unroll
    for (int a = 0; a < 7; ++a) phi[a] = 0.0f;
    return;
  }
  extrot_row_in(extrot_load(K, o), v0, v1, v2, P, phi);
}

__global__ void __launch_bounds__(256) k_extrot_rows(KeylineSoA K, const int* __restrict__ count, const float* __restrict__ vel_ptr,
                                                     int vel_stride, float* __restrict__ rowsout, const __grid_constant__ DevParams P) {
  const int b = blockIdx.y, idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count[b]) return;
  const size_t o = (size_t)b * P.kmax + idx;
  const size_t ps = (size_t)P.lanes * P.kmax;
  float phi[7];
  extrot_row(K, o, vel_ptr[(size_t)b * vel_stride], vel_ptr[(size_t)b * vel_stride + 1], vel_ptr[(size_t)b * vel_stride + 2], P, phi);
#pragma unroll
  for (int a = 0; a < 7; ++a) rowsout[a * ps + o] = phi[a];
}

// symmetric 6x6 pseudo-inverse solve standing in for TooN::SVD<6,6,float>(JtJ).backsub(JtF) (core.cpp:247-248 ->
// LAPACK sgesvd_, un-vendored): cyclic Jacobi eigen-decomposition in FP64, inverse zeroed when s_i*1e9 <= s_max
// (TooN SVD.h:37,266-274). Same rotation order and per-element arithmetic as oracle/rebvio_oracle.c:orc_solve6, executed
// by ONE WARP: lane k < 6 holds row k of A and of V in registers (all indices compile-time), a rotation (p,q) is a local
// column update, a shuffle-fed update of rows p and q, and a local V column update. Sweeps stop once the off-diagonal
// mass is below 1e-60 of the diagonal mass (FP64 converged; the oracle runs on to an exact zero, which moves X by < 1e-15).
// Call with all 32 lanes of a warp; returns x[lane] for lane < 6.
__device__ __forceinline__ double shfl_d(double v, int src) { return __shfl_sync(0xFFFFFFFFu, v, src); }
__device__ float rbv_solve6_warp(const float* A36, const float* b6) {
  const int lane = threadIdx.x & 31;
  const bool act = lane < 6;
  double a[6], v[6];
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    a[j] = act ? 0.5 * ((double)A36[lane * 6 + j] + (double)A36[j * 6 + lane]) : 0.0;
    v[j] = (act && lane == j) ? 1.0 : 0.0;
  }
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0.0, dia = 0.0;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      if (j > lane) off += a[j] * a[j];
      if (j == lane) dia += a[j] * a[j];
    }
#pragma unroll
    for (int o2 = 4; o2; o2 >>= 1) { off += __shfl_xor_sync(0xFFFFFFFFu, off, o2); dia += __shfl_xor_sync(0xFFFFFFFFu, dia, o2); }
    off = shfl_d(off, 0); dia = shfl_d(dia, 0);   // lanes 8..31 hold zeros: take lane 0's value so the exit is warp-uniform
    if (off == 0.0 || off <= 1e-60 * dia) break;
#pragma unroll
    for (int p = 0; p < 6; ++p) {
#pragma unroll
      for (int q = p + 1; q < 6; ++q) {
        const double apq = shfl_d(a[q], p);
        if (apq == 0.0) continue;
        const double app = shfl_d(a[p], p), aqq = shfl_d(a[q], q);
        const double theta = (aqq - app) / (2.0 * apq);
        const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
        const double cs = 1.0 / sqrt(t * t + 1.0), sn = t * cs;
        { const double akp = a[p], akq = a[q]; a[p] = cs * akp - sn * akq; a[q] = sn * akp + cs * akq; }   // columns p,q of every row
#pragma unroll
        for (int k = 0; k < 6; ++k) {                                                                       // rows p,q
          const double apk = shfl_d(a[k], p), aqk = shfl_d(a[k], q);
          if (lane == p) a[k] = cs * apk - sn * aqk;
          if (lane == q) a[k] = sn * apk + cs * aqk;
        }
        { const double vkp = v[p], vkq = v[q]; v[p] = cs * vkp - sn * vkq; v[q] = sn * vkp + cs * vkq; }
      }
    }
  }
  double d = 0.0;   // diagonal element of this lane's row
#pragma unroll
  for (int j = 0; j < 6; ++j) if (j == lane) d = a[j];
  double smax = act ? fabs(d) : 0.0;
#pragma unroll
  for (int o2 = 4; o2; o2 >>= 1) smax = fmax(smax, __shfl_xor_sync(0xFFFFFFFFu, smax, o2));
  smax = shfl_d(smax, 0);
  const double bk = act ? (double)b6[lane] : 0.0;
  double x = 0.0;
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    const double aii = shfl_d(d, i);
    double proj = 0.0;
#pragma unroll
    for (int k = 0; k < 6; ++k) proj += shfl_d(v[i] * bk, k);                            // proj += V[k][i] * b[k], k ascending
    if (fabs(aii) * 1e9 <= smax) continue;                                               // uniform across the warp
    proj /= aii;
    x += v[i] * proj;
  }
  return (float)x;
}

// Fast path of the same solve. The Jacobi pseudo-inverse above is a ~100-rotation dependent FP64 chain (tens of
// microseconds on one warp) that sits in the tail of extRotVel, after the last keyline block of a lane has finished. When
// J^T J is safely non-singular the pseudo-inverse IS the inverse, so: LDL^T in FP64 (every lane factorises the same 6x6 in
// registers), lane j solves for column j of the inverse, and the conditioning rule of TooN::SVD::backsub (SVD.h:266-274:
// 1/s_i is zeroed when s_i * 1e9 <= s_max) is checked rigorously without eigenvalues: s_max <= trace(A) and
// s_min >= 1 / ||A^-1||_F, so trace(A) * ||A^-1||_F * 4 < 1e9 proves that no singular value would have been dropped. Then
// x = A^-1 b equals the eigen-solve to FP64 rounding (1e-12 relative at the condition numbers seen, 44 .. 7900). Returns
// false (warp-uniform) when a pivot is not positive or the bound fails; the caller then runs the Jacobi solve.
__device__ bool rbv_solve6_ldl_warp(const float* A36, const float* b6, float& xout) {
  const int lane = threadIdx.x & 31;
  double S[6][6], L[6][6], D[6];
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 6; ++j) S[i][j] = 0.5 * ((double)A36[i * 6 + j] + (double)A36[j * 6 + i]);
  bool ok = true;
  double tr = 0.0;
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    double d = S[j][j];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k] * D[k];
    D[j] = d;
    ok = ok && (d > 0.0);
    tr += S[j][j];
    const double inv = 1.0 / d;
#pragma unroll
    for (int i = j + 1; i < 6; ++i) {
      double l = S[i][j];
#pragma unroll
      for (int k = 0; k < j; ++k) l -= L[i][k] * L[j][k] * D[k];
      L[i][j] = l * inv;
    }
  }
  const int col = lane < 6 ? lane : 0;
  double z[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) {          // L y = e_col
    double y = (i == col) ? 1.0 : 0.0;
#pragma unroll
    for (int k = 0; k < i; ++k) y -= L[i][k] * z[k];
    z[i] = y;
  }
#pragma unroll
  for (int i = 0; i < 6; ++i) z[i] /= D[i];
#pragma unroll
  for (int i = 5; i >= 0; --i)           // L^T x = z
#pragma unroll
    for (int k = i + 1; k < 6; ++k) z[i] -= L[k][i] * z[k];
  double f = 0.0, x = 0.0;
#pragma unroll
  for (int k = 0; k < 6; ++k) { f += z[k] * z[k]; x += z[k] * (double)b6[k]; }   // column col of A^-1 = row col (symmetric)
  if (lane >= 6) f = 0.0;
#pragma unroll
  for (int o2 = 4; o2; o2 >>= 1) f += __shfl_xor_sync(0xFFFFFFFFu, f, o2);
  f = shfl_d(f, 0);
  const double ninv = sqrt(f);
  ok = ok && (ninv * tr * 4.0 < 1e9);    // false for NaN / Inf as well
  xout = (float)x;
  return ok;
}

// tail of extRotVel executed by warp 0 of the reducing block: result[27] (shared) -> Wx, X, ok in lane_out
__device__ void extrot_finish_warp(const double* result, float* lane_out_b, float* s_sys /* shared, 42 floats */) {
  const int lane = threadIdx.x & 31;
  if (lane == 0) {
    int k = 0;
    for (int a = 0; a < 6; ++a)
      for (int q = a; q < 6; ++q) { s_sys[a * 6 + q] = (float)result[k]; s_sys[q * 6 + a] = (float)result[k]; ++k; }
    for (int a = 0; a < 6; ++a) s_sys[36 + a] = (float)result[21 + a];
  }
  __syncwarp();
  float x;
  if (!rbv_solve6_ldl_warp(s_sys, s_sys + 36, x)) x = rbv_solve6_warp(s_sys, s_sys + 36);   // warp-uniform branch
  for (int i = lane; i < 36; i += 32) lane_out_b[OUT_WX + i] = s_sys[i];                // _Wx = JtJ, core.cpp:249
  if (lane < 6) lane_out_b[OUT_X + lane] = x;
  const unsigned bad = __ballot_sync(0xFFFFFFFFu, lane < 6 && x != x);                  // :251-256
  if (lane == 0) lane_out_b[OUT_OK] = bad ? 0.0f : 1.0f;
}

// 27 accumulators: 21 upper-triangle entries of Phi^T Phi in (a,b) a<=b row-major order, then 6 of Phi^T Y.
__constant__ signed char c_pair_a[27] = {0,0,0,0,0,0, 1,1,1,1,1, 2,2,2,2, 3,3,3, 4,4, 5, 0,1,2,3,4,5};
__constant__ signed char c_pair_b[27] = {0,1,2,3,4,5, 1,2,3,4,5, 2,3,4,5, 3,4,5, 4,5, 5, 6,6,6,6,6,6};

__global__ void __launch_bounds__(512) k_extrot_reduce(const int* __restrict__ count, const float* __restrict__ rowsin, float* lane_out,
                                                        int ordered, const __grid_constant__ DevParams P) {
  __shared__ double sm[32 * 27];
  __shared__ double result[27];
  __shared__ float s_sys[42];
  const int b = blockIdx.x, n = count[b], t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const size_t base = (size_t)b * P.kmax;
  const size_t ps = (size_t)P.lanes * P.kmax;
  if (ordered) {
    if (warp == 0 && lane < 27) {                                                        // sequential dot, operators.hh:200-208
      const float* pa = rowsin + c_pair_a[lane] * ps + base;
      const float* pb = rowsin + c_pair_b[lane] * ps + base;
      float s = 0.0f;
      int i = 0;
      for (; i + 4 <= n; i += 4) {
        const float4 x = *reinterpret_cast<const float4*>(pa + i);
        const float4 y = *reinterpret_cast<const float4*>(pb + i);
        s += x.x * y.x; s += x.y * y.y; s += x.z * y.z; s += x.w * y.w;
      }
      for (; i < n; ++i) s += pa[i] * pb[i];
      result[lane] = (double)s;
    }
    __syncthreads();
  } else {
    double acc[27];
#pragma unroll
    for (int a = 0; a < 27; ++a) acc[a] = 0.0;
    for (int i = t; i < n; i += 512) {
      float r[7];
#pragma unroll
      for (int a = 0; a < 7; ++a) r[a] = rowsin[a * ps + base + i];
      int k = 0;
#pragma unroll
      for (int a = 0; a < 6; ++a)
#pragma unroll
        for (int q = a; q < 6; ++q) acc[k++] += (double)(r[a] * r[q]);
#pragma unroll
      for (int a = 0; a < 6; ++a) acc[21 + a] += (double)(r[a] * r[6]);
    }
    block_sum_fixed<27>(acc, sm, result);
  }
  if (warp == 0) extrot_finish_warp(result, lane_out + (size_t)b * OUT_STRIDE, s_sys);
}

#define ER_KPT 4
// RBV_REDUCE_FAST: rows + 27 sums + 6x6 solve in ONE kernel (same ticket scheme as k_tryvel_fast)
__global__ void __launch_bounds__(TV_TPB) k_extrot_fast(KeylineSoA K, const int* __restrict__ count, const float* __restrict__ vel_ptr,
                                                        int vel_stride, double* __restrict__ partials, unsigned* __restrict__ ticket,
                                                        int nblk, float* lane_out, const __grid_constant__ DevParams P) {
  __shared__ double sm[(TV_TPB / 32) * 27];
  __shared__ double result[27];
  __shared__ float s_sys[42];
  __shared__ int s_last;
  const int b = blockIdx.y, n = count[b], t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int nact = max(1, (n + TV_TPB * ER_KPT - 1) / (TV_TPB * ER_KPT));
  if ((int)blockIdx.x >= nact) return;
  // ER_KPT keylines per thread: the 27 products are accumulated in FP64 registers first, so the shuffle tree (the
  // expensive part: 27 x 5 64-bit shuffles per warp) runs once per ER_KPT x 32 keylines
  double acc[27];
#pragma unroll
  for (int a = 0; a < 27; ++a) acc[a] = 0.0;
  const float v0 = vel_ptr[(size_t)b * vel_stride], v1 = vel_ptr[(size_t)b * vel_stride + 1], v2 = vel_ptr[(size_t)b * vel_stride + 2];
  ExtRotIn in[ER_KPT];
#pragma unroll
  for (int j = 0; j < ER_KPT; ++j) {                // every load of the thread's ER_KPT keylines first (one round trip)
    const int idx = (blockIdx.x * ER_KPT + j) * TV_TPB + t;
    in[j].mid = -1;
    if (idx < n) in[j] = extrot_load(K, (size_t)b * P.kmax + idx);
  }
#pragma unroll
  for (int j = 0; j < ER_KPT; ++j) {
    if (in[j].mid < 0) continue;                    // unmatched rows are all zeros: x + 0 == x
    float r[7];
    extrot_row_in(in[j], v0, v1, v2, P, r);
    int k = 0;
#pragma unroll
    for (int a = 0; a < 6; ++a) {
#pragma unroll
      for (int q = a; q < 7; ++q) {
        const int slot = (q == 6) ? 21 + a : k++;
        acc[slot] += (double)(r[a] * r[q]);        // float product as in the reference (operators.hh:200-208), FP64 accumulate
      }
    }
  }
#pragma unroll
  for (int a = 0; a < 27; ++a) {
    double v = acc[a];
#pragma unroll
    for (int off = 16; off; off >>= 1) v += __shfl_down_sync(0xFFFFFFFFu, v, off);
    if (lane == 0) sm[warp * 27 + a] = v;
  }
  __syncthreads();
  if (t < 27) {
    double s2 = 0.0;
#pragma unroll
    for (int w = 0; w < TV_TPB / 32; ++w) s2 += sm[w * 27 + t];
    partials[((size_t)b * nblk + blockIdx.x) * 27 + t] = s2;
  }
  __threadfence();
  __syncthreads();
  if (t == 0) s_last = (atomicAdd(&ticket[b], 1u) == (unsigned)(nact - 1));
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (t < 27) {
    double s2 = 0.0;
    for (int j = 0; j < nact; ++j) s2 += __ldcg(partials + ((size_t)b * nblk + j) * 27 + t);
    result[t] = s2;
  }
  __syncthreads();
  if (t == 0) ticket[b] = 0u;
  if (warp == 0) extrot_finish_warp(result, lane_out + (size_t)b * OUT_STRIDE, s_sys);
}

// RBV_REDUCE_ORDERED: the 27 float32 products of Phi^T Phi / Phi^T Y of every keyline into the chunked addend layout (OrdER; zeros for
// unmatched keylines); k_chain_extrot behind it adds them in keyline order (TooN operators.hh:200-208) and runs the 6x6 solve.
// ord_flags != nullptr: grid (lanes, 1 + blocks), row y = 0 = one chain CTA per lane adding behind the evaluation blocks (as k_tryvel_ord*);
// nullptr: evaluation only, grid (lanes, blocks), k_chain_extrot behind it.
__global__ void __launch_bounds__(TV_TPB) k_extrot_ord(KeylineSoA K, const int* __restrict__ count, const float* __restrict__ vel_ptr, int vel_stride,
                                                       float* __restrict__ rows, int* __restrict__ ord_flags, int nblk, float* lane_out,
                                                       const __grid_constant__ DevParams P) {
  extern __shared__ __align__(128) float ord_ring[];
  __shared__ __align__(8) unsigned long long bars[2 * ORD_NST];
  __shared__ double s_res[27];
  __shared__ float s_sys[42];
  const int t = threadIdx.x;
  const int b = blockIdx.x, n = count[b];
  if (ord_flags && blockIdx.y == 0) {
    if (t >= 64) return;
    const float s = ordered_chain<OrdER>(rows + (size_t)b * OrdER::lane_stride(P.kmax), n, ord_ring, bars, nullptr, ord_flags + (size_t)b * nblk, TV_TPB * ER_KPT);
    if (t >= 32) return;
    if (t < 27) s_res[t] = (double)s;
    __syncwarp();
    extrot_finish_warp(s_res, lane_out + (size_t)b * OUT_STRIDE, s_sys);
    return;
  }
  const int blk = (int)blockIdx.y - (ord_flags ? 1 : 0);
  if (blk * TV_TPB * ER_KPT >= n) return;
  const float v0 = vel_ptr[(size_t)b * vel_stride], v1 = vel_ptr[(size_t)b * vel_stride + 1], v2 = vel_ptr[(size_t)b * vel_stride + 2];
  float* lane_rows = rows + (size_t)b * OrdER::lane_stride(P.kmax);
  ExtRotIn in[ER_KPT];
#pragma unroll
  for (int j = 0; j < ER_KPT; ++j) {                // every load of the thread's ER_KPT keylines first (one round trip)
    const int idx = (blk * ER_KPT + j) * TV_TPB + t;
    in[j].mid = -1;
    if (idx < n) in[j] = extrot_load(K, (size_t)b * P.kmax + idx);
  }
#pragma unroll
  for (int j = 0; j < ER_KPT; ++j) {
    const int idx = (blk * ER_KPT + j) * TV_TPB + t;
    if (idx >= n) continue;
    float r[7];
    extrot_row_in(in[j], v0, v1, v2, P, r);         // zeros for an unmatched keyline: x + 0 == x
    float* dst = lane_rows + OrdER::at(idx);
    int k = 0;                                      // the 27 float32 products of Phi^T Phi (upper triangle, row-major) and Phi^T Y, operators.hh:200-208
#pragma unroll
    for (int a = 0; a < 6; ++a)
#pragma unroll
      for (int q = a; q < 6; ++q) dst[(k++) * OrdER::RS] = r[a] * r[q];
#pragma unroll
    for (int a = 0; a < 6; ++a) dst[(21 + a) * OrdER::RS] = r[a] * r[6];
  }
  if (ord_flags) {   // the block's addends are out: flag for the chain CTA's loader
    __syncthreads();
    if (t == 0) {   // barrier + one cumulative fence + release store, as in tryvel_fastn_body
      __threadfence();
      asm volatile("st.release.gpu.global.s32 [%0], 1;" ::"l"(ord_flags + (size_t)b * nblk + blk) : "memory");
    }
  }
}
__global__ void __launch_bounds__(64) k_chain_extrot(const int* __restrict__ count, const float* __restrict__ rows, float* lane_out,
                                                     const __grid_constant__ DevParams P) {
  extern __shared__ __align__(128) float ord_ring[];
  __shared__ __align__(8) unsigned long long bars[2 * ORD_NST];
  __shared__ double s_res[27];
  __shared__ float s_sys[42];
  const int b = blockIdx.x, lane = threadIdx.x;
  const float s = ordered_chain<OrdER>(rows + (size_t)b * OrdER::lane_stride(P.kmax), count[b], ord_ring, bars);
  if (threadIdx.x >= 32) return;   // the loader warp
  if (lane < 27) s_res[lane] = (double)s;
  __syncwarp();
  extrot_finish_warp(s_res, lane_out + (size_t)b * OUT_STRIDE, s_sys);
}

// debug tap behind rbv_debug_read_addends: row `plane` (0..9 tryVel, 100..126 extRotVel) of one lane's addends, keylines [0, n), as the last
// evaluation kernel left them. The caller has synchronised the stream.
int tr_debug_addends(rbv_ctx* c, int plane, int lane, float* out, int n) {
  const int K = c->P.kmax;
  if (tr_ordered_legacy(c)) {
    const int a = plane >= 100 ? plane - 100 : plane;
    RBV_CUDA(cudaMemcpy(out, c->prod + ((size_t)a * c->P.lanes + lane) * K, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost));
    return RBV_OK;
  }
  const bool er = plane >= 100;
  const int a = er ? plane - 100 : plane, CH = er ? ER_CH : TV_CH;
  const size_t chunk = er ? OrdER::CHUNK : OrdTV::CHUNK, rs = er ? OrdER::RS : OrdTV::RS;
  const float* src = c->prod + (size_t)lane * (er ? OrdER::lane_stride(K) : OrdTV::lane_stride(K)) + a * rs;
  const int nch = (n + CH - 1) / CH;
  if (nch) RBV_CUDA(cudaMemcpy2D(out, sizeof(float) * CH, src, sizeof(float) * chunk, sizeof(float) * CH, (size_t)(nch - 1), cudaMemcpyDeviceToHost));
  if (nch) RBV_CUDA(cudaMemcpy(out + (size_t)(nch - 1) * CH, src + (size_t)(nch - 1) * chunk, sizeof(float) * (size_t)(n - (nch - 1) * CH), cudaMemcpyDeviceToHost));
  return RBV_OK;
}

int tr_ext_rot_vel(rbv_ctx* c, rbv_map* new_map, int vel_from_lm) {
  const DevParams& P = c->P;
  const float* vel_ptr = vel_from_lm ? reinterpret_cast<const float*>(c->lm) + offsetof(LMState, vel) / sizeof(float) : c->lane_in + IN_VEL;
  const int vel_stride = vel_from_lm ? (int)(sizeof(LMState) / sizeof(float)) : IN_STRIDE;
  if (tr_ordered_legacy(c)) {
    RBV_LAUNCH(c, k_extrot_rows, (dim3((P.keylines_max + 255) / 256, P.lanes)), (256), 0, new_map->k, new_map->count, vel_ptr, vel_stride, c->prod, P);
    RBV_LAUNCH(c, k_extrot_reduce, (P.lanes), (512), 0, new_map->count, c->prod, c->lane_out, 1, P);
  } else if (c->reduce_mode == RBV_REDUCE_ORDERED) {
    const int nb = (c->nblk + ER_KPT - 1) / ER_KPT;
    const bool two = c->tune[RBV_TUNE_ORDERED] == 2 || c->tune[RBV_TUNE_ORDERED] == 3;
    RBV_LAUNCH(c, k_extrot_ord, (dim3(P.lanes, nb + (two ? 0 : 1))), (TV_TPB), (two ? 0 : ORD_SMEM(OrdER)), new_map->k, new_map->count, vel_ptr, vel_stride,
               c->prod, two ? nullptr : c->ord_flags, c->nblk, c->lane_out, P);
    if (two) RBV_LAUNCH(c, k_chain_extrot, (P.lanes), (64), ORD_SMEM(OrdER), new_map->count, c->prod, c->lane_out, P);
  } else {
    RBV_LAUNCH(c, k_extrot_fast, (dim3((c->nblk + ER_KPT - 1) / ER_KPT, P.lanes)), (TV_TPB), 0, new_map->k, new_map->count, vel_ptr, vel_stride, c->partials, c->ticket, c->nblk, c->lane_out, P);
  }
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// directedMatch + searchMatch (edge_map.cpp:101-218): every new keyline walks its epipolar segment in the OLD map's
// pixel->id mask. Independent per keyline (the loop only mutates the querying keyline), so one thread each.
#ifndef DIR_TPB
#define DIR_TPB 128   // walks differ a lot in length: smaller blocks balance the SMs better (measured 120 -> 114 us; 64 threads: the same)
#endif
__global__ void __launch_bounds__(DIR_TPB) k_directed(KeylineSoA Kn, const int* __restrict__ count_n, KeylineSoA Ko, const int* __restrict__ mask_o,
                                                  const float* __restrict__ lane_in, int* __restrict__ lane_cnt,
                                                  unsigned long long* __restrict__ probes, const __grid_constant__ DevParams P) {
  __shared__ float s_vel[3], s_Rvel[9];
  const int b = blockIdx.y, idx = blockIdx.x * blockDim.x + threadIdx.x;
  const float* in = lane_in + (size_t)b * IN_STRIDE;
  const float* Rback = in + IN_RBACK;
  if (threadIdx.x == 0) {                                                                // :193-194
    const float* vel_in = in + IN_VEL;
    const float* Rv = in + IN_RVEL;
    float tmp[9];
    for (int r = 0; r < 3; ++r) { float acc = 0.0f; for (int j = 0; j < 3; ++j) acc += Rback[r * 3 + j] * vel_in[j]; s_vel[r] = acc; }
    for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) { float acc = 0.0f; for (int j = 0; j < 3; ++j) acc += Rback[r * 3 + j] * Rv[j * 3 + q]; tmp[r * 3 + q] = acc; }
    for (int r = 0; r < 3; ++r) for (int q = 0; q < 3; ++q) { float acc = 0.0f; for (int j = 0; j < 3; ++j) acc += tmp[r * 3 + j] * Rback[q * 3 + j]; s_Rvel[r * 3 + q] = acc; }
  }
  __syncthreads();
  int hit = -1, nprobe = 0;
  const size_t base = (size_t)b * P.kmax;
  if (idx < count_n[b]) {
    const size_t o = base + idx;
    const float fm = P.fm, pum = P.pixel_uncertainty_match, max_radius = in[IN_MAXRADIUS];
    const float2 pi = Kn.pos_img[o], g = Kn.grad[o];
    const float gn = Kn.gnorm[o], k_sr = Kn.srho[o];
    float pm[3];
#pragma unroll
    for (int r = 0; r < 3; ++r) {                                                        // :106
      float acc = 0.0f;
      acc += Rback[r * 3 + 0] * pi.x; acc += Rback[r * 3 + 1] * pi.y; acc += Rback[r * 3 + 2] * fm;
      pm[r] = acc;
    }
    const float pmx = pm[0] * fm / pm[2], pmy = pm[1] * fm / pm[2];                      // :107-108
    const float k_rho = Kn.rho[o] * fm / pm[2];                                          // :109
    const float pi0x = pmx + P.cx, pi0y = pmy + P.cy;
    float tx = -(s_vel[0] * fm - s_vel[2] * pmx), ty = -(s_vel[1] * fm - s_vel[2] * pmy);   // :114-115
    float norm_t = sqrtf(tx * tx + ty * ty);
    const float dr[3] = {fm, fm, -(pmx + pmy)};                                          // :118
    float row[3];
#pragma unroll
    for (int q = 0; q < 3; ++q) { float acc = 0.0f; for (int j = 0; j < 3; ++j) acc += dr[j] * s_Rvel[j * 3 + q]; row[q] = acc; }
    float sigma2_t = 0.0f;
#pragma unroll
    for (int j = 0; j < 3; ++j) sigma2_t += row[j] * dr[j];                              // :119
    float dq_min, dq_max, dq_rho;
    int t_steps;
    if ((double)norm_t > 1e-6) {                                                         // :126
      tx /= norm_t; ty /= norm_t;
      dq_rho = norm_t * k_rho;
      const float lo = norm_t * (k_rho - k_sr), hi = norm_t * (k_rho + k_sr);
      dq_min = (0.0f < lo ? lo : 0.0f) - pum;                                            // std::max(0, lo), :130
      dq_max = (hi < max_radius ? hi : max_radius) + pum;                                // std::min(max_radius, hi), :131
      if (dq_rho > dq_max) {
        dq_rho = (float)(0.5 * (double)(dq_max + dq_min));                               // :133
        t_steps = rbv_d2i((double)dq_rho + 0.5);
      } else {
        const float a = dq_max - dq_rho, c2 = dq_rho - dq_min;
        t_steps = rbv_f2i(a < c2 ? c2 : a);                                              // :136
      }
    } else {                                                                             // :138-149
      tx = g.x; ty = g.y;
      norm_t = gn;
      tx /= norm_t; ty /= norm_t;
      norm_t = 1.0f;
      dq_min = -max_radius - pum;
      dq_max = max_radius + pum;
      dq_rho = 0.0f;
      t_steps = rbv_f2i(dq_max);
    }
    float tn = dq_rho, tp = dq_rho + 1.0f;
    const int* m = mask_o + (size_t)b * P.npix;
    for (int ti = 0; ti < t_steps && hit < 0; ++ti, tp += 1.0f, tn -= 1.0f) {
#pragma unroll
      for (int side = 0; side < 2; ++side) {
        if (hit >= 0) break;
        float t;
        if (side) { t = tp; if (t > dq_max) continue; }
        else { t = tn; if (t < dq_min) continue; }
        const int pix = rbv_pixel_index(P.rows, P.cols, ty * t + pi0y, tx * t + pi0x);   // :164
        if (pix < 0) continue;
        ++nprobe;
        const int id = m[pix];
        if (id < 0) continue;
        const size_t oo = base + id;
        const float2 og = Ko.grad[oo];
        const float ogn = Ko.gnorm[oo];
        const float cang = (og.x * g.x + og.y * g.y) / (ogn * gn);                       // :171
        if (cang < P.cang_min_edge || fabs((double)(ogn / gn) - 1.0) > (double)P.match_threshold_norm) continue;   // :172
        const float osr = Ko.srho[oo], orho = Ko.rho[oo];
        const float v_rho_dr = (pum * pum + osr * osr * norm_t * norm_t + sigma2_t * orho * orho);   // :174
        if ((t - norm_t * orho) * (t - norm_t * orho) > v_rho_dr) continue;              // :175
        hit = id;
      }
    }
    if (hit >= 0) {                                                                      // :202-210
      const size_t oo = base + hit;
      Kn.rho[o] = Ko.rho[oo];
      Kn.srho[o] = Ko.srho[oo];
      Kn.match_id[o] = hit;
      Kn.matches[o] = Ko.matches[oo] + 1u;
      Kn.match_pos_img[o] = Ko.pos_img[oo];
      Kn.match_grad[o] = Ko.grad[oo];
      Kn.match_gnorm[o] = Ko.gnorm[oo];
      Kn.match_kf[o] = Ko.match_kf[oo];
    }
  }
  const unsigned full = 0xFFFFFFFFu;
  const int kf = (hit >= 0 && Ko.match_kf[base + hit] >= 0) ? 1 : 0;                     // :212
  const unsigned mhit = __ballot_sync(full, hit >= 0), mkf = __ballot_sync(full, kf);
#pragma unroll
  for (int off = 16; off; off >>= 1) nprobe += __shfl_down_sync(full, nprobe, off);
  if ((threadIdx.x & 31) == 0) {
    if (mhit) atomicAdd(&lane_cnt[b * CNT_STRIDE + CNT_DIR], __popc(mhit));
    if (mkf) atomicAdd(&lane_cnt[b * CNT_STRIDE + CNT_KF], __popc(mkf));
    if (nprobe) atomicAdd(&probes[b], (unsigned long long)nprobe);
  }
}

int tr_directed_match(rbv_ctx* c, rbv_map* new_map, rbv_map* old_map) {
  const DevParams& P = c->P;
  if (!c->seq_fused) RBV_LAUNCH(c, k_zero_cnt, (1), (P.lanes), 0, c->lane_cnt, CNT_DIR, c->probes);
  RBV_LAUNCH(c, k_directed, (dim3((P.keylines_max + DIR_TPB - 1) / DIR_TPB, P.lanes)), (DIR_TPB), 0, new_map->k, new_map->count, old_map->k, old_map->mask, c->lane_in, c->lane_cnt, c->probes, P);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// regularize1Iter (edge_map.cpp:220-259): two phases like the reference (evaluate everything, then write).
__global__ void __launch_bounds__(256) k_reg_eval(KeylineSoA K, const int* __restrict__ count, float* __restrict__ reg_r, float* __restrict__ reg_s,
                                                  int* __restrict__ reg_set, int* lane_cnt, int min_matches, const __grid_constant__ DevParams P) {
  const int b = blockIdx.y, idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (min_matches > 0 && lane_cnt[b * CNT_STRIDE + CNT_DIR] < min_matches) return;   // rebvio.cpp:247-252: this lane stopped (whole block)
  int set = 0;
  const size_t base = (size_t)b * P.kmax;
  if (idx < count[b]) {
    const size_t o = base + idx;
    const int in = K.id_next[o], ip = K.id_prev[o];
    if (in >= 0 && ip >= 0) {
      const size_t on = base + in, op = base + ip;
      const float rn = K.rho[on], rp = K.rho[op], sn = K.srho[on], sp = K.srho[op];
      if (!((rn - rp) * (rn - rp) > (sn * sn + sp * sp))) {                              // :235
        const float2 gn = K.grad[on], gp = K.grad[op];
        float alpha = (gn.x * gp.x + gn.y * gp.y) / (K.gnorm[on] * K.gnorm[op]);         // :238
        const float thr = P.regularization_threshold;
        if (!(alpha < thr)) {                                                            // :239
          alpha = (float)(((double)(alpha - thr)) / (1.0 - (double)thr));                // :241
          alpha = (float)((double)alpha / ((double)(fabsf(rn - rp) / (sn + sp)) + 1.0)); // :242
          const float r0 = K.rho[o], s0 = K.srho[o];
          const float wr = (float)(1.0 / (double)(s0 * s0));                             // :243
          const float wrn = alpha / (sn * sn);
          const float wrp = alpha / (sp * sp);
          reg_r[o] = (r0 * wr + rn * wrn + rp * wrp) / (wr + wrn + wrp);                 // :247-248
          reg_s[o] = (s0 * wr + sn * wrn + sp * wrp) / (wr + wrn + wrp);
          set = 1;
        }
      }
    }
    reg_set[o] = set;
  }
  const unsigned m = __ballot_sync(0xFFFFFFFFu, set);
  if ((threadIdx.x & 31) == 0 && m) atomicAdd(&lane_cnt[b * CNT_STRIDE + CNT_REG], __popc(m));
}

__global__ void k_reg_apply(KeylineSoA K, const int* __restrict__ count, const float* __restrict__ reg_r, const float* __restrict__ reg_s,
                            const int* __restrict__ reg_set, const int* __restrict__ lane_cnt, int min_matches, int kmax) {
  const int b = blockIdx.y, idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count[b]) return;
  if (min_matches > 0 && lane_cnt[b * CNT_STRIDE + CNT_DIR] < min_matches) return;
  const size_t o = (size_t)b * kmax + idx;
  if (reg_set[o]) { K.rho[o] = reg_r[o]; K.srho[o] = reg_s[o]; }                         // :252-257
}

int tr_regularize(rbv_ctx* c, rbv_map* map) {
  const DevParams& P = c->P;
  dim3 g((P.keylines_max + 255) / 256, P.lanes);
  if (!c->seq_fused) RBV_LAUNCH(c, k_zero_cnt, (1), (P.lanes), 0, c->lane_cnt, CNT_REG, nullptr);
  RBV_LAUNCH(c, k_reg_eval, (g), (256), 0, map->k, map->count, c->reg_r, c->reg_s, c->fwd_next, c->lane_cnt, c->seq_fused ? c->seq_min_matches : 0, P);
  if (c->seq_fused && map == c->field_map) c->seq_reg_pending = 1;   // k_kalman (next in rbv_track_g2, same map) applies the result
  else RBV_LAUNCH(c, k_reg_apply, (g), (256), 0, map->k, map->count, c->reg_r, c->reg_s, c->fwd_next, c->lane_cnt, c->seq_fused ? c->seq_min_matches : 0, P.kmax);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// updateInverseDepth + updateInverseDepthARLU (core.cpp:417-456): scalar EKF per matched keyline of the field map
__global__ void __launch_bounds__(256) k_kalman(KeylineSoA K, const int* __restrict__ count, const float* __restrict__ lane_in,
                                                const int* __restrict__ lane_cnt, int min_matches, const float* __restrict__ reg_r,
                                                const float* __restrict__ reg_s, const int* __restrict__ reg_set,
                                                const __grid_constant__ DevParams P) {
  const int b = blockIdx.y, idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count[b]) return;
  if (min_matches > 0 && lane_cnt[b * CNT_STRIDE + CNT_DIR] < min_matches) return;
  const size_t o = (size_t)b * P.kmax + idx;
  // every load of the keyline up front, unconditionally (all in range; unused values cost a few bytes, a dependent round trip costs a microsecond)
  float rho = K.rho[o], srho = K.srho[o];
  const int mid = K.match_id[o];
  const int rset = reg_set ? reg_set[o] : 0;
  const float rr = reg_set ? reg_r[o] : 0.0f, rs = reg_set ? reg_s[o] : 0.0f;
  const float2 q = K.pos_img[o], q0 = K.match_pos_img[o], mg = K.match_grad[o];
  const float mgn = K.match_gnorm[o];
  bool dirty = false;
  if (rset) { rho = rr; srho = rs; dirty = true; }                                       // second phase of regularize1Iter (edge_map.cpp:252-257)
  if (mid < 0) {
    if (dirty) { K.rho[o] = rho; K.srho[o] = srho; }
    return;
  }
  const float* vel = lane_in + (size_t)b * IN_STRIDE + IN_VEL;
  const float v0 = vel[0], v1 = vel[1], v2 = vel[2], fm = P.fm;
  float v_rho = srho * srho;                                                             // :429
  const float ux = mg.x / mgn, uy = mg.y / mgn;
  const float Y = ux * (q.x - q0.x) + uy * (q.y - q0.y);                                 // :432
  const float H = ux * (v0 * fm - v2 * q0.x) + uy * (v1 * fm - v2 * q0.y);               // :433
  const float rho_p = (float)(1.0 / (1.0 / (double)rho + (double)v2));                   // :434
  float F = (float)(1.0 / (1.0 + (double)(rho * v2)));                                   // :435
  F *= F;
  const float qa = P.reshape_q_abs, pu = P.pixel_uncertainty;
  const float p_p = F * v_rho * F + qa * qa;                                             // :437
  const float e = Y - H * rho_p;
  const float S = H * p_p * H + pu * pu;                                                 // :440
  const float Kg = (float)((double)(p_p * H) * (1.0 / (double)S));                       // :441
  rho = rho_p + Kg * e;                                                                  // :443
  v_rho = (float)((1.0 - (double)(Kg * H)) * (double)p_p);                               // :444
  srho = sqrtf(v_rho);
  if (rho < RBV_RHO_MIN) {                                                               // :446-455
    srho += RBV_RHO_MIN - rho;
    rho = RBV_RHO_MIN;
  } else if (rho > RBV_RHO_MAX) {
    rho = RBV_RHO_MAX;
  } else if (isnan(rho) || isnan(srho) || isinf(rho) || isinf(srho)) {
    rho = RBV_RHO_INIT;
    srho = RBV_RHO_MAX;
  }
  K.rho[o] = rho;
  K.srho[o] = srho;
}

int tr_kalman(rbv_ctx* c, rbv_map* map) {
  const DevParams& P = c->P;
  // inside rbv_track_g2 the write phase of regularize1Iter rides along (tr_regularize skipped k_reg_apply): every keyline is
  // touched by exactly one thread here and k_reg_eval has completed, so the reference's two-phase rule holds
  const bool apply = c->seq_fused && c->seq_reg_pending;
  c->seq_reg_pending = 0;
  RBV_LAUNCH(c, k_kalman, (dim3((P.keylines_max + 255) / 256, P.lanes)), (256), 0, map->k, map->count, c->lane_in, c->lane_cnt,
             c->seq_fused ? c->seq_min_matches : 0, apply ? c->reg_r : nullptr, apply ? c->reg_s : nullptr,
             apply ? (const int*)c->fwd_next : nullptr, P);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

int tr_init_ordered(void) {
  RBV_CUDA(cudaFuncSetAttribute(k_chain_tryvel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ORD_SMEM(OrdTV)));
  RBV_CUDA(cudaFuncSetAttribute(k_tryvel_ord2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ORD_SMEM(OrdTV)));
  RBV_CUDA(cudaFuncSetAttribute(k_minimize_ord2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ORD_SMEM(OrdTV)));
  RBV_CUDA(cudaFuncSetAttribute(k_minimize_ord4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ORD_SMEM(OrdTV)));
  RBV_CUDA(cudaFuncSetAttribute(k_tryvel_ord4, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ORD_SMEM(OrdTV)));
  RBV_CUDA(cudaFuncSetAttribute(k_chain_extrot, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ORD_SMEM(OrdER)));
  RBV_CUDA(cudaFuncSetAttribute(k_extrot_ord, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ORD_SMEM(OrdER)));
  return RBV_OK;
}
