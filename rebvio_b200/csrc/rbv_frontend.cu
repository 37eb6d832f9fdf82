// rbv_frontend.cu — edge front end on sm_100a: box-cascade scale space, DoG / gradient tests, sub-pixel plane fit,
// ordered keyline compaction, neighbour linking, auto threshold.
//
// Replaces (reference paths relative to the reference root):
//   FastGaussian::createIntegralImage + average   rebvio/src/scale_space.cpp:48-128   -> k_rowscan + k_colbox
//   ScaleSpace::calculateDoG / GradientMagnitude   rebvio/src/scale_space.cpp:210-233  -> fused into k_detect_flags
//   EdgeDetector::buildEdgeMap                      rebvio/src/edge_detector.cpp:45-123 -> k_detect_flags/k_detect_scan/k_detect_emit
//   EdgeDetector::joinEdges / nextKeylineIdx        rebvio/src/edge_detector.cpp:125-165 -> k_join
//   EdgeDetector::tuneThreshold                     rebvio/src/edge_detector.cpp:167-186 -> k_tune
//
// Bit-exactness: the integral images are sequential float32 prefix sums (row pass then column pass); a parallel scan
// rounds differently, so every row / column chain is walked in order by one thread and parallelism comes from the
// (lanes x rows) resp. (lanes x columns) independent chains. Compiled with -fmad=false.
#include "rbv_internal.cuh"
#include <mutex>

// ---------------------------------------------------------------------------------------------------------------
// Row pass: out[r][c] = out[r][c-1] + in[r][c]   (scale_space.cpp:50-57)
// One warp owns 32 rows; 32x32 tiles go through shared memory so that global loads/stores are 128-byte coalesced
// row segments while each lane walks its own row. The next tile is prefetched into registers while the current one
// is scanned.
__global__ void __launch_bounds__(32) k_rowscan(const float* __restrict__ in, float* __restrict__ out, int rows, int cols,
                                                size_t in_plane_stride, size_t out_plane_stride) {
  __shared__ float tile[32][33];
  const int lane = threadIdx.x;
  const int r0 = blockIdx.x * 32;
  const float* src = in + (size_t)blockIdx.y * in_plane_stride;
  float* dst = out + (size_t)blockIdx.y * out_plane_stride;
  const int nr = min(32, rows - r0);
  float nxt[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) nxt[i] = (i < nr && lane < cols) ? src[(size_t)(r0 + i) * cols + lane] : 0.0f;
  float carry = 0.0f;
  for (int c0 = 0; c0 < cols; c0 += 32) {
#pragma unroll
    for (int i = 0; i < 32; ++i) tile[i][lane] = nxt[i];
    __syncwarp();
    const int cn = c0 + 32;
    if (cn < cols) {
#pragma unroll
      for (int i = 0; i < 32; ++i) nxt[i] = (i < nr && cn + lane < cols) ? src[(size_t)(r0 + i) * cols + cn + lane] : 0.0f;
    }
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = tile[lane][j];
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      carry = (c0 + j == 0) ? v[j] : carry + v[j];
      v[j] = carry;
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) tile[lane][j] = v[j];
    __syncwarp();
    if (c0 + lane < cols) {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (i < nr) dst[(size_t)(r0 + i) * cols + c0 + lane] = tile[i][lane];
    }
    __syncwarp();
  }
}

// Row pass, 16-byte version (cols % 4 == 0, 16-byte aligned planes): the same walk with a quarter of the memory
// instructions. A warp-wide LDG.128 fetches 4 rows x 32 columns; the tile pitch of 36 floats keeps both the row-major
// float4 stores of the load phase and the lane-per-row float4 reads of the scan phase free of bank conflicts.
__global__ void __launch_bounds__(32) k_rowscan4(const float* __restrict__ in, float* __restrict__ out, int rows, int cols,
                                                 size_t in_plane_stride, size_t out_plane_stride) {
  __shared__ __align__(16) float tile[32][36];
  const int lane = threadIdx.x;
  const int r0 = blockIdx.x * 32;
  const float* src = in + (size_t)blockIdx.y * in_plane_stride;
  float* dst = out + (size_t)blockIdx.y * out_plane_stride;
  const int nr = min(32, rows - r0);
  const int lr = lane >> 3, lc = (lane & 7) * 4;     // this lane's row (+ 4k) and column offset inside a tile
  float4 nxt[8];
#pragma unroll
  for (int k = 0; k < 8; ++k)
    nxt[k] = (lr + 4 * k < nr && lc < cols) ? *reinterpret_cast<const float4*>(src + (size_t)(r0 + lr + 4 * k) * cols + lc)
                                            : make_float4(0.f, 0.f, 0.f, 0.f);
  float carry = 0.0f;
  for (int c0 = 0; c0 < cols; c0 += 32) {
#pragma unroll
    for (int k = 0; k < 8; ++k) *reinterpret_cast<float4*>(&tile[lr + 4 * k][lc]) = nxt[k];
    __syncwarp();
    const int cn = c0 + 32;
    if (cn < cols) {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        nxt[k] = (lr + 4 * k < nr && cn + lc < cols) ? *reinterpret_cast<const float4*>(src + (size_t)(r0 + lr + 4 * k) * cols + cn + lc)
                                                     : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = *reinterpret_cast<const float4*>(&tile[lane][4 * j]);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      v[j].x = (c0 + j == 0) ? v[j].x : carry + v[j].x;                                // scale_space.cpp:50-57
      v[j].y = v[j].x + v[j].y;
      v[j].z = v[j].y + v[j].z;
      v[j].w = v[j].z + v[j].w;
      carry = v[j].w;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) *reinterpret_cast<float4*>(&tile[lane][4 * j]) = v[j];
    __syncwarp();
    if (c0 + lc < cols) {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (lr + 4 * k < nr)
          *reinterpret_cast<float4*>(dst + (size_t)(r0 + lr + 4 * k) * cols + c0 + lc) = *reinterpret_cast<const float4*>(&tile[lr + 4 * k][lc]);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Column pass fused with the box average (scale_space.cpp:59-65 + 69-128).
// Thread t of a block walks integral-image column (c0 - h - 1 + t) top to bottom: ii[r][c] = ii[r-1][c] + rowsum[r][c].
// The last 32 rows of ii live in a shared-memory ring; after every chunk of CH rows the block emits the output rows whose
// lower tap row is now available. The full integral image never goes to HBM.
//   A = ii[min(r+h,R-1)][min(c+h,C-1)]  B = ii[.][c-h-1]  Cc = ii[r-h-1][.]  D = ii[r-h-1][c-h-1]
//   value = ((A-B)-Cc)+D, bottom rows ((A-Cc)-B)+D (scale_space.cpp:99 vs :119,:124), missing taps dropped,
//   times float(1.0/double(window area)) (precomputeDivisors, :130-171; the interior constant `a` (:72) is the same number).
#define CB_TPB 128
#define CB_CH 8
#define CB_RING 32   // >= 2*CB_CH + 2*hmax + 2: a chunk's writes never land on rows the previous chunk still reads -> ONE barrier per chunk

// one output value (scale_space.cpp:69-128): ro = output row, h = half width, jhi / jlo = ring columns of the right / left tap
__device__ __forceinline__ float box_value(const float (*ring)[CB_TPB], const float (*inv_area)[16], int ro, int h, int rows, int jhi,
                                           int jlo, bool left, int ncol, float inv_mid) {
  const int rlo = ro - h - 1;
  if (rlo >= 0 && ro + h < rows) {                                                     // interior rows: full window height
    const float* hi = ring[(ro + h) & (CB_RING - 1)];
    const float* lo = ring[rlo & (CB_RING - 1)];
    const float v = left ? ((hi[jhi] - hi[jlo]) - lo[jhi]) + lo[jlo] : hi[jhi] - lo[jhi];   // :99 / :94
    return v * inv_mid;
  }
  const int rhi = min(ro + h, rows - 1);
  const float* hi = ring[rhi & (CB_RING - 1)];
  float v;
  if (rlo < 0) {
    v = left ? hi[jhi] - hi[jlo] : hi[jhi];                                             // top rows :78-84
  } else {
    const float* lo = ring[rlo & (CB_RING - 1)];
    v = left ? ((hi[jhi] - lo[jhi]) - hi[jlo]) + lo[jlo] : hi[jhi] - lo[jhi];           // bottom rows :119,:124 / :114
  }
  return v * inv_area[rhi - max(rlo, -1)][ncol];
}

// ---------------------------------------------------------------------------------------------------------------
// Fused box pass (createIntegralImage + average in ONE kernel, the integral image and the row prefix never reach HBM):
// one block per image plane, so a pass moves exactly "read N + write N" floats per plane. Warp roles:
//   warp 0 (one lane)  producer: TMA bulk copies (cp.async.bulk, one per image row) of the next BP_R-row chunk into a
//                      double-buffered shared tile, completion on an mbarrier;
//   warp 1             row pass: lane = row of the chunk, in-place sequential prefix along the row (scale_space.cpp:50-57);
//   warps 2..          column pass + box: thread = image column, running column sum (:59-65) into a BP_RING-row shared ring
//                      of the integral image, then the four-tap box (:69-128) for every row whose taps are complete.
// full -> scanned -> empty mbarriers hand the chunk buffers round; the column warps use a named barrier for the ring.
// Needs cols % 4 == 0 (16-byte bulk copies) and 2*BP_R*(cols+pad) + BP_RP*roundup(cols,64) floats of smem (cols <= 768).
#define BP_R 16
#define BP_RING 32
#define BP_CH 8

// (mbarrier / TMA bulk-copy helpers: rbv_internal.cuh)

// The ring of the integral image is COLUMN-major: ring[col * BP_RP + (row & 31)], rows 32..39 of a column mirror rows 0..7.
// With the row index in the low-order position, the 8 rows of a sub-chunk are 8 consecutive words behind one base register
// (immediate offsets, no wrap: start <= 31, start + 7 <= 38), so the steady state spends no instruction on tap addresses; the
// odd column pitch keeps lane = column accesses conflict-free.
#define BP_RP 41
// one output value (border rows); jhi / jlo are image columns
__device__ __forceinline__ float box_value_fw(const float* __restrict__ ring, const float (*inv_area)[16], int ro, int h, int rows,
                                              int jhi, int jlo, bool left, int ncol, float inv_mid) {
  const int rlo = ro - h - 1;
  const float* chi = ring + jhi * BP_RP;
  const float* clo = ring + max(jlo, 0) * BP_RP;
  if (rlo >= 0 && ro + h < rows) {
    const int rh = (ro + h) & (BP_RING - 1), rl = rlo & (BP_RING - 1);
    const float v = left ? ((chi[rh] - clo[rh]) - chi[rl]) + clo[rl] : chi[rh] - chi[rl];   // :99 / :94
    return v * inv_mid;
  }
  const int rhi = min(ro + h, rows - 1);
  const int rh = rhi & (BP_RING - 1);
  float v;
  if (rlo < 0) {
    v = left ? chi[rh] - clo[rh] : chi[rh];                                             // :78-84
  } else {
    const int rl = rlo & (BP_RING - 1);
    v = left ? ((chi[rh] - chi[rl]) - clo[rh]) + clo[rl] : chi[rh] - chi[rl];           // :119,:124 / :114
  }
  return v * inv_area[rhi - max(rlo, -1)][ncol];
}

// shared-memory accesses by 32-bit shared address: keeps the address arithmetic of the hot loops in 32-bit integer adds
__device__ __forceinline__ float lds_f32(unsigned a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));   // volatile: ordered against the barrier / store asms, but ALU work and global stores may move across it
  return v;
}
__device__ __forceinline__ float2 lds_f32x2(unsigned a) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts_f32(unsigned a, float x) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(x) : "memory"); }
__device__ __forceinline__ void sts_f32x2(unsigned a, float x, float y) {
  asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a), "f"(x), "f"(y) : "memory");
}

// taps of one output column: shared addresses of ring row 0 of the right / left tap column, divisor, left-tap flag
struct BpTap {
  unsigned p_hi, p_lo;
  float inv;
  bool left;
};
__device__ __forceinline__ BpTap bp_tap(unsigned ring_a, int c, int h, int cols) {
  BpTap t;
  const int jhi = min(c + h, cols - 1), jlo = c - h - 1;
  t.left = jlo >= 0;
  t.p_hi = ring_a + 4u * BP_RP * (unsigned)jhi;
  t.p_lo = ring_a + 4u * BP_RP * (unsigned)max(jlo, 0);
  t.inv = (float)(1.0 / (double)((float)(2 * h + 1) * (float)(jhi - max(jlo, -1))));     // interior-row divisor (:72 / :168)
  return t;
}

// BP_CH interior output rows of two columns (t0: column c, t1: column c + 32; dst = output row ro0, column c). Output i takes
// ring rows (ro0 + h + i) and (ro0 - h - 1 + i): ((A - B) - C) + D (:99), or A - C where there is no left tap (:94).
template <bool MIXED>
__device__ __forceinline__ void bp_box_rows(const BpTap& t0, const BpTap& t1, int ro0, int h, float* __restrict__ dst, bool ok0, bool ok1,
                                            int cols) {
  const unsigned hs = 4u * (unsigned)((ro0 + h) & (BP_RING - 1)), ls = 4u * (unsigned)((ro0 - h - 1) & (BP_RING - 1));
  const unsigned pa0 = t0.p_hi + hs, pb0 = t0.p_lo + hs, pc0 = t0.p_hi + ls, pd0 = t0.p_lo + ls;
  const unsigned pa1 = t1.p_hi + hs, pb1 = t1.p_lo + hs, pc1 = t1.p_hi + ls, pd1 = t1.p_lo + ls;
#pragma unroll
  for (int i = 0; i < BP_CH; ++i) {
    const float a0 = lds_f32(pa0 + 4u * i), b0 = lds_f32(pb0 + 4u * i), c0 = lds_f32(pc0 + 4u * i), d0 = lds_f32(pd0 + 4u * i);
    const float a1 = lds_f32(pa1 + 4u * i), b1 = lds_f32(pb1 + 4u * i), c1 = lds_f32(pc1 + 4u * i), d1 = lds_f32(pd1 + 4u * i);
    float v0 = ((a0 - b0) - c0) + d0, v1 = ((a1 - b1) - c1) + d1;
    if (MIXED) {
      if (!t0.left) v0 = a0 - c0;
      if (!t1.left) v1 = a1 - c1;
    }
    if (ok0) dst[0] = v0 * t0.inv;
    if (ok1) dst[32] = v1 * t1.inv;
    dst += cols;
  }
}

__device__ __forceinline__ void bp_scan4(float4& v, float& carry, bool first) {
  v.x = first ? v.x : carry + v.x;
  v.y = v.x + v.y;
  v.z = v.y + v.z;
  v.w = v.z + v.w;
  carry = v.w;
}

// ---------------------------------------------------------------------------------------------------------------
// Column pass + box of the split path (few planes: single stream, or images too wide for k_boxpass).
// dual = 1 (first pass): both scales are box-filtered from ONE column scan of the shared row-prefix image (plane =
// blockIdx.y = lane; outputs out0 / out1 with widths d0 / d1). dual = 0: plane = blockIdx.y, width d0 or d1 by plane % nz.
// With one warp per scheduler this kernel lives or dies by instruction latency, so the steady state is straight-line
// code: CB_CH rows are scanned, one barrier, then CB_CH output rows whose 4 x CB_CH taps are all issued before the first
// subtraction; the row-prefix loads run two chunks ahead.
template <bool MIXED>
__device__ __forceinline__ void cb_box_rows(unsigned p_hi, unsigned p_lo, bool left, float inv, int ro0, int h, float* __restrict__ dst, int cols) {
  constexpr unsigned PITCH = CB_TPB * 4u, WRAP = CB_RING * PITCH - 1u;
  unsigned hi_off = (unsigned)((ro0 + h) & (CB_RING - 1)) * PITCH;
  unsigned lo_off = (unsigned)((ro0 - h - 1) & (CB_RING - 1)) * PITCH;
  float a[CB_CH], bq[CB_CH], cq[CB_CH], dq[CB_CH];
#pragma unroll
  for (int i = 0; i < CB_CH; ++i) {
    a[i] = lds_f32(p_hi + hi_off); bq[i] = lds_f32(p_lo + hi_off);
    cq[i] = lds_f32(p_hi + lo_off); dq[i] = lds_f32(p_lo + lo_off);
    hi_off = (hi_off + PITCH) & WRAP;
    lo_off = (lo_off + PITCH) & WRAP;
  }
#pragma unroll
  for (int i = 0; i < CB_CH; ++i) {
    float v = ((a[i] - bq[i]) - cq[i]) + dq[i];                                          // :99
    if (MIXED && !left) v = a[i] - cq[i];                                              // :94
    dst[(size_t)i * cols] = v * inv;
  }
}

__global__ void __launch_bounds__(CB_TPB) k_colbox(const float* __restrict__ rowsum, float* __restrict__ out0, float* __restrict__ out1,
                                                   int rows, int cols, size_t in_plane_stride, size_t out_plane_stride, int nz, int d0,
                                                   int d1, int dual) {
  __shared__ float ring[CB_RING][CB_TPB];
  __shared__ float inv_area[16][16];
  const int t = threadIdx.x;
  const int z = dual ? 0 : blockIdx.y % nz;
  const int ha = (z ? d1 : d0) / 2;                 // half width of output A
  const int hb = dual ? d1 / 2 : ha;                // half width of output B (dual only)
  const int hm = max(ha, hb);
  const int W = CB_TPB - (2 * hm + 1);
  const int c0 = blockIdx.x * W;
  const int col = c0 - hm - 1 + t;                  // integral-image column scanned by this thread
  const bool scan_ok = col >= 0 && col < cols;
  const int oc = c0 + t;                            // output column of this thread (t < W)
  const bool out_ok = t < W && oc < cols;
  const float* src = rowsum + (size_t)blockIdx.y * in_plane_stride + (scan_ok ? col : 0);
  float* dstA = out0 + (size_t)blockIdx.y * out_plane_stride;
  float* dstB = out1 + (size_t)blockIdx.y * out_plane_stride;
  for (int i = t; i < 256; i += CB_TPB) {
    int a = i / 16, b = i % 16;
    inv_area[a][b] = (a && b) ? (float)(1.0 / (double)((float)a * (float)b)) : 0.0f;   // precomputeDivisors :130-171 (== `a` of :72 inside)
  }
  // taps of output A / B in ring coordinates (ring column j <-> image column c0 - hm - 1 + j)
  const int chiA = min(oc + ha, cols - 1), cloA = oc - ha - 1;
  const int jhiA = out_ok ? chiA - (c0 - hm - 1) : 0, jloA = t + (hm - ha), ncolA = chiA - max(cloA, -1);
  const int chiB = min(oc + hb, cols - 1), cloB = oc - hb - 1;
  const int jhiB = out_ok ? chiB - (c0 - hm - 1) : 0, jloB = t + (hm - hb), ncolB = chiB - max(cloB, -1);
  const bool leftA = cloA >= 0, leftB = cloB >= 0;
  const float invA = out_ok ? (float)(1.0 / (double)((float)(2 * ha + 1) * (float)ncolA)) : 0.0f;
  const float invB = out_ok ? (float)(1.0 / (double)((float)(2 * hb + 1) * (float)ncolB)) : 0.0f;
  const unsigned ring_a = smem_u32(&ring[0][0]);
  const unsigned pA_hi = ring_a + 4u * (unsigned)jhiA, pA_lo = ring_a + 4u * (unsigned)(leftA ? jloA : 0);
  const unsigned pB_hi = ring_a + 4u * (unsigned)jhiB, pB_lo = ring_a + 4u * (unsigned)(leftB ? jloB : 0);
  const bool mixed = blockIdx.x == 0;               // only the first strip has columns without a left tap
  float run = 0.0f;
  float n0[CB_CH], n1[CB_CH];
#pragma unroll
  for (int i = 0; i < CB_CH; ++i) {
    n0[i] = (scan_ok && i < rows) ? src[(size_t)i * cols] : 0.0f;
    n1[i] = (scan_ok && CB_CH + i < rows) ? src[(size_t)(CB_CH + i) * cols] : 0.0f;
  }
  int onext = 0;
  for (int k0 = 0; k0 < rows; k0 += CB_CH) {
    const int k1 = min(k0 + CB_CH, rows);
    float cur[CB_CH];
#pragma unroll
    for (int i = 0; i < CB_CH; ++i) { cur[i] = n0[i]; n0[i] = n1[i]; }
    const int kn = k0 + 2 * CB_CH;
    if (kn + CB_CH <= rows) {
      if (scan_ok) {
#pragma unroll
        for (int i = 0; i < CB_CH; ++i) n1[i] = src[(size_t)(kn + i) * cols];
      }
    } else {
#pragma unroll
      for (int i = 0; i < CB_CH; ++i) n1[i] = (scan_ok && kn + i < rows) ? src[(size_t)(kn + i) * cols] : 0.0f;
    }
    const bool fast = k0 >= 2 * CB_CH && k0 + CB_CH < rows;
    if (fast) {
      const unsigned ra = ring_a + ((unsigned)(k0 & (CB_RING - 1)) * CB_TPB + (unsigned)t) * 4u;   // k0 % CB_CH == 0: no wrap inside
#pragma unroll
      for (int i = 0; i < CB_CH; ++i) { run = run + cur[i]; sts_f32(ra + i * CB_TPB * 4u, run); }   // column pass :59-65
    } else {
#pragma unroll
      for (int i = 0; i < CB_CH; ++i) {
        if (k0 + i < rows) {
          run = (k0 + i == 0) ? cur[i] : run + cur[i];
          ring[(k0 + i) & (CB_RING - 1)][t] = run;
        }
      }
    }
    __syncthreads();
    const int olast = (k1 == rows) ? rows - 1 : k1 - 1 - hm;   // last output row whose taps are all in the ring
    if (out_ok) {
      if (fast) {   // onext == k0 - hm
        const int ro0 = k0 - hm;
        float* dA = dstA + (size_t)ro0 * cols + oc;
        if (mixed) cb_box_rows<true>(pA_hi, pA_lo, leftA, invA, ro0, ha, dA, cols);
        else cb_box_rows<false>(pA_hi, pA_lo, leftA, invA, ro0, ha, dA, cols);
        if (dual) {
          float* dB = dstB + (size_t)ro0 * cols + oc;
          if (mixed) cb_box_rows<true>(pB_hi, pB_lo, leftB, invB, ro0, hb, dB, cols);
          else cb_box_rows<false>(pB_hi, pB_lo, leftB, invB, ro0, hb, dB, cols);
        }
      } else {
        for (int ro = onext; ro <= olast; ++ro) {
          dstA[(size_t)ro * cols + oc] = box_value(ring, inv_area, ro, ha, rows, jhiA, jloA, leftA, ncolA, invA);
          if (dual) dstB[(size_t)ro * cols + oc] = box_value(ring, inv_area, ro, hb, rows, jhiB, jloB, leftB, ncolB, invB);
        }
      }
    }
    onext = max(onext, olast + 1);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Latency-optimised box pass for FEW planes (single stream). With one or two planes the split kernels above are bound by
// global-load latency: a warp fetches one 32-column tile per round trip, 24 round trips per row block (k_rowscan4 22 us,
// k_colbox 41 us per launch at 752x480) although the dependent add chains themselves are only 752 resp. 480 adds long
// (~1.5 us). Here a block brings its WHOLE working set into shared memory with one round trip - 32 full rows by TMA bulk
// copies (row pass), a 32-column strip of all rows by 4-byte cp.async (column pass) - then walks the chains out of shared
// memory, and every thread of the block shares the load, box and store work. Same arithmetic, statement by statement.
#define RL_R 32
#define RL_TPB 256
__global__ void __launch_bounds__(RL_TPB) k_rowscan_lat(const float* __restrict__ in, float* __restrict__ out, int rows, int cols,
                                                        size_t in_plane_stride, size_t out_plane_stride, int spitch, int rpb) {
  extern __shared__ __align__(128) float rl_smem[];   // [rpb][spitch] (rpb <= 32 rows per block), (spitch / 4) odd: conflict-free 128-bit row walks
  __shared__ __align__(8) unsigned long long rl_bar;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int r0 = blockIdx.x * rpb, nr = min(rpb, rows - r0);
  const float* src = in + (size_t)blockIdx.y * in_plane_stride + (size_t)r0 * cols;
  float* dst = out + (size_t)blockIdx.y * out_plane_stride + (size_t)r0 * cols;
  const int n4 = cols >> 2;
  if (tid == 0) {
    mbar_init(&rl_bar, 1u);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (warp == 0) {
    if (lane == 0) mbar_arrive_expect_tx(&rl_bar, (unsigned)(nr * cols * 4));
    __syncwarp();
    if (lane < nr) bulk_g2s(rl_smem + (size_t)lane * spitch, src + (size_t)lane * cols, (unsigned)(cols * 4), &rl_bar);
    mbar_wait(&rl_bar, 0u);
    if (lane < nr) {                                                                     // scale_space.cpp:50-57, lane = row
      float4* row = reinterpret_cast<float4*>(rl_smem + (size_t)lane * spitch);
      float carry = 0.0f;
      float4 a[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = (i < n4) ? row[i] : make_float4(0.f, 0.f, 0.f, 0.f);
      for (int c4 = 0; c4 < n4; c4 += 4) {
        float4 nx[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) nx[i] = (c4 + 4 + i < n4) ? row[c4 + 4 + i] : make_float4(0.f, 0.f, 0.f, 0.f);
        bp_scan4(a[0], carry, c4 == 0);
        bp_scan4(a[1], carry, false);
        bp_scan4(a[2], carry, false);
        bp_scan4(a[3], carry, false);
#pragma unroll
        for (int i = 0; i < 4; ++i) if (c4 + i < n4) row[c4 + i] = a[i];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = nx[i];
      }
    }
  }
  __syncthreads();
  for (int r = warp; r < nr; r += RL_TPB / 32) {
    const float4* srow = reinterpret_cast<const float4*>(rl_smem + (size_t)r * spitch);
    float4* drow = reinterpret_cast<float4*>(dst + (size_t)r * cols);
    for (int c4 = lane; c4 < n4; c4 += 32) drow[c4] = srow[c4];
  }
}

// one output value from a linear [rows][tw] tile of the integral image (same regions and association orders as box_value)
__device__ __forceinline__ float box_value_lin(const float* __restrict__ tile, int tw, const float (*inv_area)[16], int ro, int h, int rows,
                                               int jhi, int jlo, bool left, int ncol, float inv_mid) {
  const int rlo = ro - h - 1;
  if (rlo >= 0 && ro + h < rows) {
    const float* hi = tile + (ro + h) * tw;
    const float* lo = tile + rlo * tw;
    const float v = left ? ((hi[jhi] - hi[jlo]) - lo[jhi]) + lo[jlo] : hi[jhi] - lo[jhi];   // :99 / :94
    return v * inv_mid;
  }
  const int rhi = min(ro + h, rows - 1);
  const float* hi = tile + rhi * tw;
  float v;
  if (rlo < 0) {
    v = left ? hi[jhi] - hi[jlo] : hi[jhi];                                             // :78-84
  } else {
    const float* lo = tile + rlo * tw;
    v = left ? ((hi[jhi] - lo[jhi]) - hi[jlo]) + lo[jlo] : hi[jhi] - lo[jhi];           // :119,:124 / :114
  }
  return v * inv_area[rhi - max(rlo, -1)][ncol];
}

#define CL_W 32
#define CL_TPB 256
__global__ void __launch_bounds__(CL_TPB) k_colbox_lat(const float* __restrict__ rowsum, float* __restrict__ out0, float* __restrict__ out1,
                                                       int rows, int cols, size_t in_plane_stride, size_t out_plane_stride, int nz, int d0,
                                                       int d1, int dual, int hmax, int cw) {
  extern __shared__ __align__(16) float cl_tile[];   // [rows][tw]: image columns c0 - hmax - 1 .. c0 + cw - 1 + hmax (cw <= 32, even)
  __shared__ float inv_area[16][16];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int plane = blockIdx.y;
  const int z = dual ? 0 : plane % nz;
  const int ha = (z ? d1 : d0) / 2;
  const int hb = dual ? d1 / 2 : ha;
  const int tw = cw + 2 * hmax + 1;                   // odd: lane = column accesses are conflict-free
  const int c0 = blockIdx.x * cw, cbase = c0 - hmax - 1;
  const float* src = rowsum + (size_t)plane * in_plane_stride;
  for (int i = tid; i < 256; i += CL_TPB) {
    const int a = i / 16, b = i % 16;
    inv_area[a][b] = (a && b) ? (float)(1.0 / (double)((float)a * (float)b)) : 0.0f;   // precomputeDivisors :130-171
  }
  {  // the whole strip in one round trip: 4-byte cp.async straight into shared memory, all of a thread's copies in flight
    const int total = rows * tw, q = CL_TPB / tw, rem = CL_TPB % tw;
    int r = tid / tw, j = tid % tw;
    const unsigned tile_a = smem_u32(cl_tile);
    for (int i = tid; i < total; i += CL_TPB) {
      const int col = cbase + j;
      if (col >= 0 && col < cols)
        asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(tile_a + 4u * (unsigned)(r * tw + j)), "l"(src + (size_t)r * cols + col) : "memory");
      r += q; j += rem;
      if (j >= tw) { j -= tw; ++r; }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  }
  __syncthreads();
  if (tid < tw) {                                      // column pass :59-65, thread = column of the strip
    const int col = cbase + tid;
    if (col >= 0 && col < cols) {
      float* p = cl_tile + tid;
      float run = p[0];
      int r = 1;
      for (; r + 8 <= rows; r += 8) {
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = p[(r + i) * tw];
#pragma unroll
        for (int i = 0; i < 8; ++i) { run = run + v[i]; p[(r + i) * tw] = run; }
      }
      for (; r < rows; ++r) { run = run + p[r * tw]; p[r * tw] = run; }
    }
  }
  __syncthreads();
  const int col = c0 + lane;
  if (lane >= cw || col >= cols) return;
  int jhi[2], jlo[2], ncol[2];
  float inv_mid[2];
  const int hh[2] = {ha, hb};
#pragma unroll
  for (int w = 0; w < 2; ++w) {
    const int chi = min(col + hh[w], cols - 1), clo = col - hh[w] - 1;
    jhi[w] = chi - cbase; jlo[w] = max(clo, 0) - cbase; ncol[w] = chi - max(clo, -1);
    inv_mid[w] = (float)(1.0 / (double)((float)(2 * hh[w] + 1) * (float)ncol[w]));     // interior-row divisor (:72 / :168)
  }
  const bool leftA = col - ha - 1 >= 0, leftB = col - hb - 1 >= 0;
  float* dA = out0 + (size_t)plane * out_plane_stride + col;
  float* dB = out1 + (size_t)plane * out_plane_stride + col;
  for (int ro = warp; ro < rows; ro += CL_TPB / 32) {
    dA[(size_t)ro * cols] = box_value_lin(cl_tile, tw, inv_area, ro, ha, rows, jhi[0], jlo[0], leftA, ncol[0], inv_mid[0]);
    if (dual) dB[(size_t)ro * cols] = box_value_lin(cl_tile, tw, inv_area, ro, hb, rows, jhi[1], jlo[1], leftB, ncol[1], inv_mid[1]);
  }
}

#define BP_MAX_THREADS 640
// column warps among `nwarps` warps: all but warp 0 (producer), warp 1 (row pass) and warps 5, 9, 13, ... (left idle)
__host__ __device__ inline int bp_ncw(int nwarps) {
  int n = 0;
  for (int w = 2; w < nwarps; ++w) n += (w & 3) != 1;
  return n;
}

__global__ void __launch_bounds__(BP_MAX_THREADS, 1) k_boxpass(const float* __restrict__ in, float* __restrict__ out0, float* __restrict__ out1,
                                                     int rows, int cols, size_t in_plane_stride, size_t out_plane_stride, int nz, int d0,
                                                     int d1, int dual, int spitch, int in_div) {
  extern __shared__ __align__(128) float bp_smem[];
  __shared__ float inv_area[16][16];
  __shared__ __align__(8) unsigned long long bars[6];   // full[2], scanned[2], empty[2]
  float* S = bp_smem;                                  // [2][BP_R][spitch]
  float* ring = bp_smem + 2 * BP_R * spitch;           // [columns][BP_RP], see box_value_fw
  unsigned long long* full = bars;
  unsigned long long* scanned = bars + 2;
  unsigned long long* empty = bars + 4;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // warp roles: 0 producer, 1 row pass; warps 5, 9, 13, ... exit so that the row-pass warp has its scheduler (warp id % 4)
  // to itself (its dependent add chain is the critical path); every other warp is a column warp.
  const int ncw = bp_ncw((int)(blockDim.x >> 5));
  const int plane = blockIdx.x;
  const int z = dual ? 0 : plane % nz;
  const int ha = (z ? d1 : d0) / 2;
  const int hb = dual ? d1 / 2 : ha;
  const int hm = max(ha, hb);
  const float* src = in + (size_t)(plane / in_div) * in_plane_stride;   // in_div = 2: both scales of a lane read the same input plane
  for (int i = tid; i < 256; i += blockDim.x) {
    const int a = i / 16, b = i % 16;
    inv_area[a][b] = (a && b) ? (float)(1.0 / (double)((float)a * (float)b)) : 0.0f;   // precomputeDivisors :130-171
  }
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(full + i, 1u); mbar_init(scanned + i, 32u); mbar_init(empty + i, (unsigned)ncw); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int nchunks = (rows + BP_R - 1) / BP_R;
  if (warp == 0) {
    // lane 0 waits for the buffer and arms the barrier; lanes 0..nr-1 then issue one row copy each (a single thread
    // needs ~125 cycles per UBLKCP, which would put 2000 cycles of issue time on every chunk)
    for (int ci = 0; ci < nchunks; ++ci) {
      const int buf = ci & 1;
      const int nr = min(BP_R, rows - ci * BP_R);
      if (lane == 0) {
        if (ci >= 2) mbar_wait(empty + buf, (unsigned)(((ci >> 1) - 1) & 1));
        // (the consumers fence generic -> async proxy before they arrive on `empty`)
        mbar_arrive_expect_tx(full + buf, (unsigned)(nr * cols * 4));
      }
      __syncwarp();
      if (lane < nr)
        bulk_g2s(S + (size_t)(buf * BP_R + lane) * spitch, src + (size_t)(ci * BP_R + lane) * cols, (unsigned)(cols * 4), full + buf);
    }
  } else if (warp == 1) {
    const int n4 = cols >> 2;
    for (int ci = 0; ci < nchunks; ++ci) {
      const int buf = ci & 1;
      mbar_wait(full + buf, (unsigned)((ci >> 1) & 1));
      if (lane < BP_R) {
        float4* row = reinterpret_cast<float4*>(S + (size_t)(buf * BP_R + lane) * spitch);
        float carry = 0.0f;
        float4 a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = (i < n4) ? row[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        for (int c4 = 0; c4 < n4; c4 += 4) {
          float4 nx[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) nx[i] = (c4 + 4 + i < n4) ? row[c4 + 4 + i] : make_float4(0.f, 0.f, 0.f, 0.f);
          bp_scan4(a[0], carry, c4 == 0);
          bp_scan4(a[1], carry, false);
          bp_scan4(a[2], carry, false);
          bp_scan4(a[3], carry, false);
#pragma unroll
          for (int i = 0; i < 4; ++i) if (c4 + i < n4) row[c4 + i] = a[i];
#pragma unroll
          for (int i = 0; i < 4; ++i) a[i] = nx[i];
        }
      }
      mbar_arrive(scanned + buf);
    }
  } else if ((warp & 3) != 1) {
    const int q = (warp - 2) - ((warp - 2) >> 2);       // column-warp index
    // this thread owns columns c0 and c0 + 32: consecutive lanes = consecutive columns, so every shared access of the warp
    // is conflict-free (row-major S: unit stride; column-major ring: stride BP_RP, odd) and every global store a full
    // 128-byte segment. Columns >= cols (last warp) run on in-range garbage: S and the ring are padded to ncw * 64 columns,
    // their results are never stored to global memory and no valid column taps them (jhi is clamped to cols - 1).
    const int c0 = q * 64 + lane, c1 = c0 + 32;
    const bool okc[2] = {c0 < cols, c1 < cols};
    const int cc[2] = {min(c0, cols - 1), min(c1, cols - 1)};   // tap columns of the (unused) out-of-range outputs stay in range
    float* dstA = out0 + (size_t)plane * out_plane_stride;
    float* dstB = out1 + (size_t)plane * out_plane_stride;
    const unsigned ring_a = smem_u32(ring), S_a = smem_u32(S);
    const unsigned spb = 4u * (unsigned)spitch;
    const unsigned rcol = ring_a + 4u * BP_RP * (unsigned)c0;        // ring column c0, row 0 (column c1: + 32 * BP_RP words)
    constexpr unsigned RC1 = 4u * BP_RP * 32u;
    const BpTap tA0 = bp_tap(ring_a, cc[0], ha, cols), tA1 = bp_tap(ring_a, cc[1], ha, cols);
    const BpTap tB0 = bp_tap(ring_a, cc[0], hb, cols), tB1 = bp_tap(ring_a, cc[1], hb, cols);
    const bool mixed = q == 0;                          // only the first few columns lack a left tap
    // generic (border-row) path
    int jhi[2][2], jlo[2][2], ncol[2][2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      jhi[0][j] = min(cc[j] + ha, cols - 1); jlo[0][j] = cc[j] - ha - 1; ncol[0][j] = jhi[0][j] - max(jlo[0][j], -1);
      jhi[1][j] = min(cc[j] + hb, cols - 1); jlo[1][j] = cc[j] - hb - 1; ncol[1][j] = jhi[1][j] - max(jlo[1][j], -1);
    }
    const unsigned nbar = (unsigned)(ncw * 32);
    float run0 = 0.0f, run1 = 0.0f;
    int onext = 0;
    for (int ci = 0; ci < nchunks; ++ci) {
      const int buf = ci & 1;
      if (q == 0) mbar_wait(scanned + buf, (unsigned)((ci >> 1) & 1));   // one warp watches the mbarrier,
      asm volatile("bar.sync 1, %0;" ::"r"(nbar) : "memory");                  // the others sleep on the hardware barrier
      const int kend = min((ci + 1) * BP_R, rows);
#pragma unroll
      for (int sub = 0; sub < BP_R / BP_CH; ++sub) {
        const int k0 = ci * BP_R + sub * BP_CH;
        if (k0 >= rows) break;
        const int k1 = min(k0 + BP_CH, rows);
        // steady state: a full sub-chunk whose BP_CH output rows (k0 - hm ...) are interior in the row direction
        const bool fast = k0 >= 2 * BP_CH && k0 + BP_CH < rows;
        {
          const unsigned sa = S_a + (unsigned)((buf * BP_R + sub * BP_CH) * spitch + c0) * 4u;
          const int kr = k0 & (BP_RING - 1);                     // k0 % BP_CH == 0: no wrap inside the sub-chunk
          const unsigned ra = rcol + 4u * (unsigned)kr;
          if (fast) {
            float v0[BP_CH], v1[BP_CH];
#pragma unroll
            for (int i = 0; i < BP_CH; ++i) { v0[i] = lds_f32(sa + i * spb); v1[i] = lds_f32(sa + i * spb + 128u); }
#pragma unroll
            for (int i = 0; i < BP_CH; ++i) {                                            // column pass :59-65
              run0 = run0 + v0[i]; run1 = run1 + v1[i];
              v0[i] = run0; v1[i] = run1;
              sts_f32(ra + 4u * i, run0);
              sts_f32(ra + RC1 + 4u * i, run1);
            }
            if (kr == 0) {                                                               // rows 0..7 again as rows 32..39
#pragma unroll
              for (int i = 0; i < BP_CH; ++i) {
                sts_f32(ra + 4u * (BP_RING + i), v0[i]);
                sts_f32(ra + RC1 + 4u * (BP_RING + i), v1[i]);
              }
            }
          } else {
#pragma unroll
            for (int i = 0; i < BP_CH; ++i) {
              if (k0 + i < rows) {
                const float x0 = lds_f32(sa + i * spb), x1 = lds_f32(sa + i * spb + 128u);
                run0 = (k0 + i == 0) ? x0 : run0 + x0;
                run1 = (k0 + i == 0) ? x1 : run1 + x1;
                sts_f32(ra + 4u * i, run0);
                sts_f32(ra + RC1 + 4u * i, run1);
                if (kr == 0) {
                  sts_f32(ra + 4u * (BP_RING + i), run0);
                  sts_f32(ra + RC1 + 4u * (BP_RING + i), run1);
                }
              }
            }
          }
        }
        if (k1 == kend) {   // last read of this chunk buffer: hand it back to the producer
          __syncwarp();
          if (lane == 0) { fence_proxy_async(); mbar_arrive(empty + buf); }
        }
        asm volatile("bar.sync 1, %0;" ::"r"(nbar) : "memory");
        const int olast = (k1 == rows) ? rows - 1 : k1 - 1 - hm;
        if (fast) {   // onext == k0 - hm here
          const int ro0 = k0 - hm;
          float* dA = dstA + (size_t)ro0 * cols + c0;
          if (mixed) bp_box_rows<true>(tA0, tA1, ro0, ha, dA, okc[0], okc[1], cols);
          else bp_box_rows<false>(tA0, tA1, ro0, ha, dA, okc[0], okc[1], cols);
          if (dual) {
            float* dB = dstB + (size_t)ro0 * cols + c0;
            if (mixed) bp_box_rows<true>(tB0, tB1, ro0, hb, dB, okc[0], okc[1], cols);
            else bp_box_rows<false>(tB0, tB1, ro0, hb, dB, okc[0], okc[1], cols);
          }
        } else {
          for (int ro = onext; ro <= olast; ++ro) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              if (!okc[j]) continue;
              const int cj = j ? c1 : c0;
              dstA[(size_t)ro * cols + cj] =
                  box_value_fw(ring, inv_area, ro, ha, rows, jhi[0][j], jlo[0][j], jlo[0][j] >= 0, ncol[0][j], j ? tA1.inv : tA0.inv);
              if (dual)
                dstB[(size_t)ro * cols + cj] =
                    box_value_fw(ring, inv_area, ro, hb, rows, jhi[1][j], jlo[1][j], jlo[1][j] >= 0, ncol[1][j], j ? tB1.inv : tB0.inv);
            }
          }
        }
        onext = max(onext, olast + 1);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Fused box pass, second form (k_boxpass2). Same roles as k_boxpass (TMA producer, row pass, column pass + box; one CTA per
// plane, neither the row prefix nor the integral image reaches HBM) and the same arithmetic statement by statement, but the
// box no longer reads its four taps from a 40-row shared ring (k_boxpass: 6 shared-memory accesses per pixel, `mio_throttle`
// its top stall). The thread that scans integral-image column j now produces output column j - h:
//   A = ii[r+h][j]      is the running column sum it has just formed,
//   C = ii[r-h-1][j]    is the sum it formed W = 2h+1 rows earlier: a W-deep delay line in registers,
//   B = ii[r+h][j-W]    is read from the ring (the one tap load left), and kept in a second delay line, because
//   D = ii[r-h-1][j-W]  is the B of W rows earlier.
// Rows are walked in sub-chunks of W rows, so both delay lines are register arrays with compile-time indices (two sets that
// swap roles every sub-chunk: no copies). Per pixel: 1 load of the row prefix, 1 store + 1 load of the ring = 3 shared accesses, and
// the ring shrinks to 2 x W rows, which pays for a THIRD TMA chunk buffer (fill / row pass / column pass of three chunks overlap)
// and for a two-stage row pass (warp RA scans the left half of a chunk's rows and hands the 16 carries to warp RB, which scans
// the right half while RA starts the next chunk: the 752-add dependent chain per chunk becomes two pipelined chains of 376).
// The last h output columns (right tap clamped to cols-1), the top rows (no upper taps) and the bottom h rows (lower taps clamped,
// other association order, scale_space.cpp:114-124) take generic paths that read all their taps from the ring.
// Box widths 7 and 9 (the reference's sigma: {7,7,7} / {9,9,9}); other widths fall back to k_boxpass.
#define B2_R 16
#define B2_NB 3
#define B2_SROWS (B2_R * B2_NB)
#define B2_MAX_THREADS 512

template <int W> struct B2Cols { float v0[W], v1[W], b0[W], b1[W]; };   // column sums and left-tap values of one sub-chunk, columns c0 and c0 + 32

__device__ __forceinline__ void b2_scan_range(float4* row, int b4, int e4, float& carry) {   // scale_space.cpp:50-57 over float4 [b4, e4)
  float4 a[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) a[i] = (b4 + i < e4) ? row[b4 + i] : make_float4(0.f, 0.f, 0.f, 0.f);
  for (int c4 = b4; c4 < e4; c4 += 4) {
    float4 nx[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) nx[i] = (c4 + 4 + i < e4) ? row[c4 + 4 + i] : make_float4(0.f, 0.f, 0.f, 0.f);
    bp_scan4(a[0], carry, c4 == 0);
    bp_scan4(a[1], carry, false);
    bp_scan4(a[2], carry, false);
    bp_scan4(a[3], carry, false);
#pragma unroll
    for (int i = 0; i < 4; ++i) if (c4 + i < e4) row[c4 + i] = a[i];
#pragma unroll
    for (int i = 0; i < 4; ++i) a[i] = nx[i];
  }
}

// what a column thread keeps across sub-chunks
struct B2State {
  unsigned S_a, ring_a, spb, rpb;     // shared addresses of S[0][c0] and ring[0][0][c0]; byte pitches of an S row and a ring row
  unsigned nbar;
  int rows, cols, nchunks, c0, h, q, lane;
  float run0, run1, inv0, inv1;
  bool st0, st1;
  int ci_ready, ci_rel;
  float* dst;                         // output plane + (c0 - h)
  unsigned long long *scanned, *empty;
  const float (*inv_area)[16];
};

// one sub-chunk: rows k0 .. k0+nr-1 are scanned into `cur` (and the ring), then the box rows k0-h .. k0+nr-1-h are written from
// `cur` (A, B) and `prev` (C, D). TOP = first sub-chunk (no upper taps, scale_space.cpp:78-84). Column warp 0 also holds the columns
// without a left tap (:94) and computes the right-border columns; that is a run-time test so that all column warps share one copy of the
// code (the first version of this kernel had eight copies of this body - two W x warp 0 / others x ping-pong - 11 k instructions,
// and spent a quarter of its samples waiting for instruction fetch).
template <int W, bool TOP>
__device__ __forceinline__ void b2_sub(B2State& st, B2Cols<W>& cur, const B2Cols<W>& prev, int s) {
  constexpr int h = W / 2;
  const int k0 = s * W, nr = min(W, st.rows - k0), k1 = k0 + nr;
  const int need = (k1 - 1) / B2_R;
  if (st.q == 0)
    for (; st.ci_ready <= need; ++st.ci_ready) mbar_wait(st.scanned + st.ci_ready % B2_NB, (unsigned)((st.ci_ready / B2_NB) & 1));
  asm volatile("bar.sync 1, %0;" ::"r"(st.nbar) : "memory");   // chunk rows visible to every column warp; the box reads of sub-chunk s-1 are done
  const unsigned rb = st.ring_a + (unsigned)((s & 1) * W) * st.rpb;
  {
    int srow = k0 % B2_SROWS;
    float x0[W], x1[W];
#pragma unroll
    for (int i = 0; i < W; ++i) {
      const unsigned sa = st.S_a + (unsigned)srow * st.spb;
      x0[i] = lds_f32(sa); x1[i] = lds_f32(sa + 128u);
      srow = (srow + 1 == B2_SROWS) ? 0 : srow + 1;
    }
#pragma unroll
    for (int i = 0; i < W; ++i) {                                                       // column pass, scale_space.cpp:59-65
      if (i < nr) {
        if (TOP && i == 0) { st.run0 = x0[0]; st.run1 = x1[0]; }
        else { st.run0 = st.run0 + x0[i]; st.run1 = st.run1 + x1[i]; }
        cur.v0[i] = st.run0; cur.v1[i] = st.run1;
        sts_f32(rb + (unsigned)i * st.rpb, st.run0);
        sts_f32(rb + (unsigned)i * st.rpb + 128u, st.run1);
      }
    }
  }
  // chunk buffers whose last row has now been read go back to the producer
  while (st.ci_rel < st.nchunks && ((st.ci_rel + 1) * B2_R <= k1 || k1 == st.rows)) {
    __syncwarp();
    if (st.lane == 0) { fence_proxy_async(); mbar_arrive(st.empty + st.ci_rel % B2_NB); }
    ++st.ci_rel;
  }
  asm volatile("bar.sync 1, %0;" ::"r"(st.nbar) : "memory");   // ring rows of this sub-chunk complete
  const bool left0 = st.c0 >= W;
  {
    const unsigned lb = rb - 4u * (unsigned)(left0 ? W : 0);   // column c0 - W (clamped to c0 where there is no left tap: never used)
#pragma unroll
    for (int i = 0; i < W; ++i) {
      if (i < nr) { cur.b0[i] = lds_f32(lb + (unsigned)i * st.rpb); cur.b1[i] = lds_f32(rb - 4u * W + (unsigned)i * st.rpb + 128u); }
    }
  }
  float* d = st.dst + (ptrdiff_t)(k0 - h) * st.cols;
#pragma unroll
  for (int i = 0; i < W; ++i) {
    if (i < nr && (!TOP || i >= h)) {
      float o0, o1;
      if (TOP) {                                                                        // :78-84: rows without upper taps
        o0 = left0 ? cur.v0[i] - cur.b0[i] : cur.v0[i];
        o1 = cur.v1[i] - cur.b1[i];
        const int nc0 = min(W, st.c0 + 1);
        o0 *= st.inv_area[i + 1][nc0];
        o1 *= st.inv_area[i + 1][W];
      } else {                                                                          // :99 ((A - B) - C) + D, :94 A - C
        o0 = ((cur.v0[i] - cur.b0[i]) - prev.v0[i]) + prev.b0[i];
        if (!left0) o0 = cur.v0[i] - prev.v0[i];
        o1 = ((cur.v1[i] - cur.b1[i]) - prev.v1[i]) + prev.b1[i];
        o0 *= st.inv0; o1 *= st.inv1;
      }
      if (st.st0) d[(size_t)i * st.cols] = o0;
      if (st.st1) d[(size_t)i * st.cols + 32] = o1;
    }
  }
  // the h output columns whose right tap is clamped to the last column: every tap from the ring (this sub-chunk and the previous one);
  // the (row, column) pairs are dealt out over the lanes of column warp 0
  if (st.q == 0) {
    const unsigned ring0 = st.ring_a - 4u * (unsigned)st.c0;   // column 0 of ring row 0
    const unsigned ca = ring0 + (unsigned)((s & 1) * W) * st.rpb, pa = ring0 + (unsigned)(((s & 1) ^ 1) * W) * st.rpb;
    for (int idx = st.lane; idx < h * nr; idx += 32) {
      const int i = idx / h, j = idx - i * h;
      if (TOP && i < h) continue;
      const int oc = st.cols - h + j, jlo = oc - h - 1, ncol = (st.cols - 1) - jlo;
      const float A = lds_f32(ca + (unsigned)i * st.rpb + 4u * (unsigned)(st.cols - 1)), Bv = lds_f32(ca + (unsigned)i * st.rpb + 4u * (unsigned)jlo);
      float o;
      if (TOP) o = (A - Bv) * st.inv_area[i + 1][ncol];
      else {
        const float C = lds_f32(pa + (unsigned)i * st.rpb + 4u * (unsigned)(st.cols - 1)), D = lds_f32(pa + (unsigned)i * st.rpb + 4u * (unsigned)jlo);
        o = (((A - Bv) - C) + D) * (float)(1.0 / (double)((float)W * (float)ncol));
      }
      st.dst[(ptrdiff_t)(k0 - h + i) * st.cols - (st.c0 - h) + oc] = o;
    }
  }
}

template <int W>
__device__ __forceinline__ void boxpass2_body(const float* __restrict__ src, float* __restrict__ outp, int rows, int cols, int spitch, int ncw, int rcols,
                                              float* bp_smem, float (*inv_area)[16], unsigned long long* bars, float (*carry_s)[B2_R]) {
  constexpr int h = W / 2;
  float* S = bp_smem;                                   // [B2_SROWS][spitch]: image row r in S row r % 48
  float* ring = bp_smem + B2_SROWS * spitch;            // [2][W][rcols]: integral-image rows of the current / previous sub-chunk
  unsigned long long* full = bars;                      // [3] TMA landed
  unsigned long long* scanned = bars + 3;               // [3] row pass done
  unsigned long long* empty = bars + 6;                 // [3] column warps are done with the chunk
  unsigned long long* half = bars + 9;                  // [2] RA has left its carries for RB
  unsigned long long* cfree = bars + 11;                // [2] RB has taken them
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 256; i += blockDim.x) {
    const int a = i / 16, b = i % 16;
    inv_area[a][b] = (a && b) ? (float)(1.0 / (double)((float)a * (float)b)) : 0.0f;   // precomputeDivisors :130-171
  }
  if (tid == 0) {
    for (int i = 0; i < B2_NB; ++i) { mbar_init(full + i, 1u); mbar_init(scanned + i, 32u); mbar_init(empty + i, (unsigned)ncw); }
    for (int i = 0; i < 2; ++i) { mbar_init(half + i, 32u); mbar_init(cfree + i, 32u); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int nchunks = (rows + B2_R - 1) / B2_R;
  const int n4 = cols >> 2;
  const int n4a = min(n4, ((n4 / 2 + 3) / 4) * 4);      // RA scans float4 [0, n4a), RB [n4a, n4)
  if (warp == 0) {
    for (int ci = 0; ci < nchunks; ++ci) {
      const int buf = ci % B2_NB;
      const int nr = min(B2_R, rows - ci * B2_R);
      if (lane == 0) {
        if (ci >= B2_NB) mbar_wait(empty + buf, (unsigned)(((ci / B2_NB) - 1) & 1));
        mbar_arrive_expect_tx(full + buf, (unsigned)(nr * cols * 4));
      }
      __syncwarp();
      if (lane < nr)
        bulk_g2s(S + (size_t)(buf * B2_R + lane) * spitch, src + (size_t)(ci * B2_R + lane) * cols, (unsigned)(cols * 4), full + buf);
    }
  } else if (warp == 1) {                               // RA: left half of every row of the chunk, carries to RB
    for (int ci = 0; ci < nchunks; ++ci) {
      const int buf = ci % B2_NB;
      const int nr = min(B2_R, rows - ci * B2_R);
      mbar_wait(full + buf, (unsigned)((ci / B2_NB) & 1));
      float carry = 0.0f;
      if (lane < nr) b2_scan_range(reinterpret_cast<float4*>(S + (size_t)(buf * B2_R + lane) * spitch), 0, n4a, carry);
      if (ci >= 2) mbar_wait(cfree + (ci & 1), (unsigned)(((ci >> 1) - 1) & 1));
      if (lane < B2_R) carry_s[ci & 1][lane] = carry;
      mbar_arrive(half + (ci & 1));
    }
  } else if (warp == 2) {                               // RB: right half, starting from RA's carries
    for (int ci = 0; ci < nchunks; ++ci) {
      const int buf = ci % B2_NB;
      const int nr = min(B2_R, rows - ci * B2_R);
      mbar_wait(full + buf, (unsigned)((ci / B2_NB) & 1));
      mbar_wait(half + (ci & 1), (unsigned)((ci >> 1) & 1));
      float carry = (lane < B2_R) ? carry_s[ci & 1][lane] : 0.0f;
      __syncwarp();
      mbar_arrive(cfree + (ci & 1));
      if (lane < nr && n4a < n4) b2_scan_range(reinterpret_cast<float4*>(S + (size_t)(buf * B2_R + lane) * spitch), n4a, n4, carry);
      mbar_arrive(scanned + buf);
    }
  } else if (warp - 3 < ncw) {
    B2State st;
    st.q = warp - 3; st.lane = lane;
    st.c0 = st.q * 64 + lane;
    const int c1 = st.c0 + 32;
    st.rows = rows; st.cols = cols; st.nchunks = nchunks; st.h = h;
    st.spb = 4u * (unsigned)spitch; st.rpb = 4u * (unsigned)rcols;
    st.S_a = smem_u32(S) + 4u * (unsigned)st.c0;
    st.ring_a = smem_u32(ring) + 4u * (unsigned)st.c0;
    st.nbar = (unsigned)(ncw * 32);
    st.run0 = st.run1 = 0.0f;
    st.inv0 = (float)(1.0 / (double)((float)W * (float)min(W, st.c0 + 1)));              // interior-row divisor (:72 / :168)
    st.inv1 = (float)(1.0 / (double)((float)W * (float)W));
    st.st0 = st.c0 >= h && st.c0 < cols; st.st1 = c1 < cols;
    st.ci_ready = 0; st.ci_rel = 0;
    st.dst = outp + (st.c0 - h);
    st.scanned = scanned; st.empty = empty; st.inv_area = inv_area;
    B2Cols<W> cur, prev;
    const int nsub = (rows + W - 1) / W;
    b2_sub<W, true>(st, cur, prev, 0);
#pragma unroll 1
    for (int s2 = 1; s2 < nsub; ++s2) {
      prev = cur;                                       // 4 W register moves against ~500 instructions of body: one copy of the code
      b2_sub<W, false>(st, cur, prev, s2);
    }
    // bottom h rows (:104-124): lower taps clamped to the last row, ((A - C) - B) + D; every tap from the ring, which holds the last two
    // sub-chunks (row r in buffer (r / W) & 1, slot r % W). No barrier needed: the ring is not written any more.
    const unsigned ring0 = st.ring_a - 4u * (unsigned)st.c0;
    auto ring_at = [&](int r, int c) -> float { return lds_f32(ring0 + (unsigned)(((r / W) & 1) * W + r % W) * st.rpb + 4u * (unsigned)c); };
    for (int ro = rows - h; ro < rows; ++ro) {
      const int rlo = ro - h - 1, rhi = rows - 1, nrow = rhi - rlo;
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        int oc;
        if (j == 0) { if (!st.st0) continue; oc = st.c0 - h; }
        else if (j == 1) { if (!st.st1) continue; oc = c1 - h; }
        else { if (!(st.q == 0 && lane < h)) continue; oc = cols - h + lane; }
        const int jhi = min(oc + h, cols - 1), jlo = oc - h - 1, ncol = jhi - max(jlo, -1);
        const float A = ring_at(rhi, jhi), C = ring_at(rlo, jhi);
        float v;
        if (jlo >= 0) { const float Bv = ring_at(rhi, jlo), D = ring_at(rlo, jlo); v = ((A - C) - Bv) + D; }   // :119, :124
        else v = A - C;                                                                                       // :114
        outp[(size_t)ro * cols + oc] = v * inv_area[nrow][ncol];
      }
    }
  }
}

__global__ void __launch_bounds__(B2_MAX_THREADS, 1) k_boxpass2(const float* __restrict__ in, float* __restrict__ out, int rows, int cols,
                                                                 size_t in_plane_stride, size_t out_plane_stride, int nz, int d0, int d1, int spitch,
                                                                 int in_div, int ncw, int rcols) {
  extern __shared__ __align__(128) float bp_smem[];
  __shared__ float inv_area[16][16];
  __shared__ __align__(8) unsigned long long bars[13];
  __shared__ float carry_s[2][B2_R];
  const int plane = blockIdx.x;
  const int d = (plane % nz) ? d1 : d0;
  const float* src = in + (size_t)(plane / in_div) * in_plane_stride;   // in_div = 2: both scales of a lane read the same input plane
  float* outp = out + (size_t)plane * out_plane_stride;
  if (d == 7) boxpass2_body<7>(src, outp, rows, cols, spitch, ncw, rcols, bp_smem, inv_area, bars, carry_s);
  else boxpass2_body<9>(src, outp, rows, cols, spitch, ncw, rcols, bp_smem, inv_area, bars, carry_s);
}

static int b2_ncw(int cols) { return (cols + 63) / 64; }
static int b2_spitch(int cols) { return cols + (((cols >> 2) & 1) ? 8 : 4); }
static size_t b2_smem_bytes(int cols) {
  return sizeof(float) * ((size_t)B2_SROWS * b2_spitch(cols) + (size_t)2 * 9 * (b2_ncw(cols) * 64) + 64);
}
static bool b2_supported(const rbv_ctx* c) {
  const int cols = c->P.cols;
  for (int s2 = 0; s2 < 2; ++s2)
    for (int p = 0; p < 3; ++p) if (c->widths[s2][p] != 7 && c->widths[s2][p] != 9) return false;
  return (cols & 3) == 0 && (3 + b2_ncw(cols)) * 32 <= B2_MAX_THREADS && b2_smem_bytes(cols) <= 225 * 1024 && c->P.rows >= 32 && cols >= 64;
}
static int b2_launch(rbv_ctx* c, const float* in, float* out, int planes, size_t in_stride, size_t out_stride, int nz, int d0, int d1, int in_div = 1) {
  const int cols = c->P.cols, ncw = b2_ncw(cols);
  RBV_LAUNCH(c, k_boxpass2, (planes), ((3 + ncw) * 32), b2_smem_bytes(cols), in, out, c->P.rows, cols, in_stride, out_stride, nz, d0, d1,
             b2_spitch(cols), in_div, ncw, ncw * 64);
  return RBV_OK;
}

static int bp_spitch(int cols) { return cols + (((cols >> 2) & 1) ? 8 : 4); }   // (pitch / 4) odd: conflict-free 128-bit row accesses
static int bp_threads(int cols) {   // smallest block whose column warps cover cols at two columns per thread
  const int need = (cols + 63) / 64;
  int w = 3;
  while (bp_ncw(w) < need) ++w;
  return w * 32;
}
// S: two chunk buffers [BP_R][spitch]; ring: [columns padded to 64][BP_RP] (+ 64 floats: out-of-range columns of the last
// warp read S up to 63 floats past a row, which for the last row of buffer 1 is the start of the ring - in range either way)
static size_t bp_smem_bytes(int cols) {
  return sizeof(float) * ((size_t)2 * BP_R * bp_spitch(cols) + (size_t)((cols + 63) / 64 * 64) * BP_RP + 64);
}
static bool bp_supported(const rbv_ctx* c) {
  const int cols = c->P.cols;
  return (cols & 3) == 0 && bp_threads(cols) <= BP_MAX_THREADS && bp_smem_bytes(cols) <= 225 * 1024;
}
// which fused kernel: k_boxpass (automatic) or k_boxpass2 (register delay lines) with RBV_TUNE_BOX = 4
static bool b2_use(const rbv_ctx* c) { return b2_supported(c) && c->tune[RBV_TUNE_BOX] == 4; }   // opt-in: measured slower than k_boxpass so far (230 us vs 105 us)
static bool bp_use(const rbv_ctx* c, const void* src) {   // a fused (one block per plane) box pass at all?
  if (((uintptr_t)src & 15) != 0) return false;
  if (c->tune[RBV_TUNE_BOX] == 1 || c->tune[RBV_TUNE_BOX] == 3) return false;
  if (c->tune[RBV_TUNE_BOX] == 2) return bp_supported(c);
  if (c->tune[RBV_TUNE_BOX] == 4) return b2_supported(c);
  return bp_supported(c) && c->P.lanes >= 16;   // one block per plane: with few planes the split kernels have far more parallelism
}
static int bp_launch(rbv_ctx* c, const float* in, float* out0, float* out1, int planes, size_t in_stride, size_t out_stride, int nz,
                     int d0, int d1, int dual, int in_div = 1) {
  const int cols = c->P.cols;
  const int threads = bp_threads(cols);
  RBV_LAUNCH(c, k_boxpass, (planes), (threads), bp_smem_bytes(cols), in, out0, out1, c->P.rows, cols, in_stride, out_stride, nz, d0, d1,
             dual, bp_spitch(cols), in_div);
  return RBV_OK;
}

// ---- latency-optimised pair (k_rowscan_lat / k_colbox_lat) ----
static int lat_spitch(int cols) { return cols + (((cols >> 2) & 1) ? 8 : 4); }
static int lat_hmax(const rbv_ctx* c) {
  int m = 0;
  for (int s2 = 0; s2 < 2; ++s2)
    for (int p = 0; p < 3; ++p) m = c->widths[s2][p] / 2 > m ? c->widths[s2][p] / 2 : m;
  return m;
}
// rows per block of the row kernel / output columns per block of the column kernel: the largest of 32, 16, 8 whose tile fits
static int lat_rpb(const rbv_ctx* c) {
  for (int r = 32; r >= 8; r >>= 1)
    if (sizeof(float) * (size_t)r * lat_spitch(c->P.cols) <= 200 * 1024) return r;
  return 0;
}
static int lat_cw(const rbv_ctx* c) {
  for (int w = CL_W; w >= 8; w >>= 1)
    if (sizeof(float) * (size_t)c->P.rows * (w + 2 * lat_hmax(c) + 1) <= 200 * 1024) return w;
  return 0;
}
static size_t lat_row_smem(const rbv_ctx* c) { return sizeof(float) * (size_t)lat_rpb(c) * lat_spitch(c->P.cols); }
static size_t lat_col_smem(const rbv_ctx* c) { return sizeof(float) * (size_t)c->P.rows * (lat_cw(c) + 2 * lat_hmax(c) + 1); }
static bool lat_use(const rbv_ctx* c, const void* src) {
  const bool fits = (c->P.cols & 3) == 0 && lat_rpb(c) > 0 && lat_cw(c) > 0 && lat_hmax(c) <= 7 && ((uintptr_t)src & 15) == 0;
  if (!fits || c->tune[RBV_TUNE_BOX] == 1 || c->tune[RBV_TUNE_BOX] == 2) return false;
  if (c->tune[RBV_TUNE_BOX] == 3) return true;
  // automatic: one scanning warp per block pays off only while the grid does not fill the GPU (few planes); with many
  // planes the split kernels hide their load latency behind other blocks and move more bytes per second
  // (and only with full-size tiles: at 4K the 8-row / 8-column tiles measured no better than the split kernels)
  return lat_rpb(c) == 32 && lat_cw(c) == CL_W && 2 * c->P.lanes * ((c->P.rows + 31) / 32) <= 2 * 148;
}
static void lat_rowscan(rbv_ctx* c, const float* in, float* out, int planes) {
  const size_t N = (size_t)c->P.npix;
  const int rpb = lat_rpb(c);
  RBV_LAUNCH(c, k_rowscan_lat, (dim3((c->P.rows + rpb - 1) / rpb, planes)), (RL_TPB), lat_row_smem(c), in, out, c->P.rows, c->P.cols, N, N,
             lat_spitch(c->P.cols), rpb);
}
static void lat_colbox(rbv_ctx* c, const float* rowsum, float* out0, float* out1, int planes, size_t in_stride, size_t out_stride, int nz,
                       int d0, int d1, int dual) {
  const int cw = lat_cw(c);
  RBV_LAUNCH(c, k_colbox_lat, (dim3((c->P.cols + cw - 1) / cw, planes)), (CL_TPB), lat_col_smem(c), rowsum, out0, out1, c->P.rows,
             c->P.cols, in_stride, out_stride, nz, d0, d1, dual, lat_hmax(c), cw);
}

int fe_init(rbv_ctx* c) {
  (void)c;   // the current device is the context's; the limit is per device and shared by every context on it
  RBV_CUDA(cudaFuncSetAttribute(k_boxpass, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
  RBV_CUDA(cudaFuncSetAttribute(k_boxpass2, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
  RBV_CUDA(cudaFuncSetAttribute(k_rowscan_lat, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  RBV_CUDA(cudaFuncSetAttribute(k_colbox_lat, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  return RBV_OK;
}

static void fe_rowscan(rbv_ctx* c, const float* in, float* out, int planes) {
  const int rows = c->P.rows, cols = c->P.cols;
  const size_t N = (size_t)c->P.npix;
  if ((cols & 3) == 0 && (((uintptr_t)in | (uintptr_t)out) & 15) == 0)
    RBV_LAUNCH(c, k_rowscan4, (dim3((rows + 31) / 32, planes)), (32), 0, in, out, rows, cols, N, N);
  else
    RBV_LAUNCH(c, k_rowscan, (dim3((rows + 31) / 32, planes)), (32), 0, in, out, rows, cols, N, N);
}

// ScaleSpace::build (scale_space.cpp:203-208): 3 box passes per scale, both scales from the same input.
// Buffers are [lane][scale][N]; result: scale s of lane b = bufA[(b*2+s)*N].
// part: 0 = everything, 1 = only the kernel that reads d_img (the one launch whose arguments change from frame to frame),
// 2 = everything after it (captured in a CUDA graph by the caller)
int fe_scale_space(rbv_ctx* c, const float* d_img, int part) {
  const int rows = c->P.rows, cols = c->P.cols, B = c->P.lanes;
  const size_t N = (size_t)c->P.npix;
  if (bp_use(c, d_img)) {   // fused passes: 64 N -> 44 N bytes of traffic per frame, 3 launches instead of 6
    // pass 0: one CTA per (lane, scale), both reading the lane's frame. The dual form (one CTA per lane filtering both
    // widths from one scan) moves a third less data but is a 64-CTA launch of 140 us on 148 SMs; two single-width CTAs per
    // lane fill the GPU like passes 1 and 2 do.
    const int pl = 2 * B;
    if (b2_use(c)) {
      if (part != 2) b2_launch(c, d_img, c->bufA, pl, N, N, 2, c->widths[0][0], c->widths[1][0], 2);
      if (part != 1) {
        b2_launch(c, c->bufA, c->bufB, pl, N, N, 2, c->widths[0][1], c->widths[1][1]);
        b2_launch(c, c->bufB, c->bufA, pl, N, N, 2, c->widths[0][2], c->widths[1][2]);
      }
    } else {
      if (part != 2) bp_launch(c, d_img, c->bufA, c->bufA, pl, N, N, 2, c->widths[0][0], c->widths[1][0], 0, 2);
      if (part != 1) {
        bp_launch(c, c->bufA, c->bufB, c->bufB, pl, N, N, 2, c->widths[0][1], c->widths[1][1], 0);
        bp_launch(c, c->bufB, c->bufA, c->bufA, pl, N, N, 2, c->widths[0][2], c->widths[1][2], 0);
      }
    }
    RBV_CUDA(cudaGetLastError());
    return RBV_OK;
  }
  if (lat_use(c, d_img)) {   // few planes: whole-working-set-in-shared-memory kernels (one load round trip per block)
    if (part != 2) lat_rowscan(c, d_img, c->rowsum, B);
    if (part == 1) { RBV_CUDA(cudaGetLastError()); return RBV_OK; }
    lat_colbox(c, c->rowsum, c->bufA, c->bufA + N, B, N, 2 * N, 1, c->widths[0][0], c->widths[1][0], 1);
    float* src = c->bufA;
    float* dst = c->bufB;
    for (int p = 1; p < 3; ++p) {
      lat_rowscan(c, src, c->rowsum, 2 * B);
      lat_colbox(c, c->rowsum, dst, dst, 2 * B, N, N, 2, c->widths[0][p], c->widths[1][p], 0);
      float* t = src; src = dst; dst = t;
    }
    RBV_CUDA(cudaGetLastError());
    return RBV_OK;
  }
  // pass 0: one row prefix per lane, one column scan feeding both box widths (widths[0][0], widths[1][0])
  if (part != 2) fe_rowscan(c, d_img, c->rowsum, B);
  if (part == 1) { RBV_CUDA(cudaGetLastError()); return RBV_OK; }
  {
    const int d0 = c->widths[0][0], d1 = c->widths[1][0];
    const int hmax = (d0 > d1 ? d0 : d1) / 2;
    const int W = CB_TPB - (2 * hmax + 1);
    RBV_LAUNCH(c, k_colbox, (dim3((cols + W - 1) / W, B)), (CB_TPB), 0, c->rowsum, c->bufA, c->bufA + N, rows, cols, N, 2 * N, 1, d0, d1, 1);
  }
  float* src = c->bufA;
  float* dst = c->bufB;
  for (int p = 1; p < 3; ++p) {
    fe_rowscan(c, src, c->rowsum, 2 * B);
    const int d0 = c->widths[0][p], d1 = c->widths[1][p];
    const int hmax = (d0 > d1 ? d0 : d1) / 2;
    const int W = CB_TPB - (2 * hmax + 1);
    RBV_LAUNCH(c, k_colbox, (dim3((cols + W - 1) / W, 2 * B)), (CB_TPB), 0, c->rowsum, dst, dst, rows, cols, N, N, 2, d0, d1, 0);
    float* t = src; src = dst; dst = t;
  }
  // after passes 1 and 2 the result is back in bufA
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}
// which of the three scale-space paths fe_scale_space takes for this input pointer (its alignment decides): 2 = fused box pass, 1 = latency
// pair, 0 = split kernels. Part of the detect graph's key: part 1 (launched directly) and the captured part 2 must be the same path.
int fe_uses_fused_box(rbv_ctx* c, const float* d_img) { return bp_use(c, d_img) ? (b2_use(c) ? 3 : 2) : (lat_use(c, d_img) ? 1 : 0); }

// debug: one box pass (createIntegralImage + average) on lane 0, plane 0 of the scratch buffers
int fe_box_pass_debug(rbv_ctx* c, int scale, int pass, const float* d_in, float* d_out) {
  const int rows = c->P.rows, cols = c->P.cols;
  const size_t N = (size_t)c->P.npix;
  const int d = c->widths[scale][pass];
  if (c->tune[RBV_TUNE_BOX] == 4 && bp_use(c, d_in) && b2_use(c)) {
    b2_launch(c, d_in, d_out, 1, N, N, 1, d, d);
    RBV_CUDA(cudaGetLastError());
    return RBV_OK;
  }
  if (c->tune[RBV_TUNE_BOX] == 2 && bp_use(c, d_in)) {
    bp_launch(c, d_in, d_out, d_out, 1, N, N, 1, d, d, 0);
    RBV_CUDA(cudaGetLastError());
    return RBV_OK;
  }
  if (c->tune[RBV_TUNE_BOX] == 3 && lat_use(c, d_in)) {
    lat_rowscan(c, d_in, c->rowsum, 1);
    lat_colbox(c, c->rowsum, d_out, d_out, 1, N, N, 1, d, d, 0);
    RBV_CUDA(cudaGetLastError());
    return RBV_OK;
  }
  fe_rowscan(c, d_in, c->rowsum, 1);
  const int W = CB_TPB - (2 * (d / 2) + 1);
  RBV_LAUNCH(c, k_colbox, (dim3((cols + W - 1) / W, 1)), (CB_TPB), 0, c->rowsum, d_out, d_out, rows, cols, N, N, 1, d, d, 0);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Per-pixel keyline test (edge_detector.cpp:83-107). y[25] = DoG window in row-major order (dr outer, dc inner).
struct FitResult {
  float th0, th1, xs, ys;
};
template <bool PN_DONE = false>
__device__ __forceinline__ bool rbv_fit_plane(const float* y, const DevParams& P, float gradient_threshold_squared, FitResult& o) {
  if (!PN_DONE) {
    int pn = 0;
#pragma unroll
    for (int k = 0; k < 25; ++k) pn = (y[k] > 0.0f) ? pn + 1 : pn - 1;                 // :92
    if (fabs((double)pn) > (double)P.pn_threshold) return false;                       // :96
  }
  float th[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {                                                        // :98 sequential dot from 0
    float acc = 0.0f;
#pragma unroll
    for (int k = 0; k < 25; ++k) acc += P.pinv[a * 25 + k] * y[k];
    th[a] = acc;
  }
  const float n2 = th[0] * th[0] + th[1] * th[1];
  const float tmp = th[2] / n2;                                                        // :99
  const float xs = -th[0] * tmp, ys = -th[1] * tmp;
  if (fabsf(xs) > 0.5f || fabsf(ys) > 0.5f) return false;                              // :103
  if (n2 < gradient_threshold_squared) return false;                                   // :107
  o.th0 = th[0]; o.th1 = th[1]; o.xs = xs; o.ys = ys;
  return true;
}

// Pass 1: flags. Block = 256 threads, tile = 16 rows x 64 columns (+2 halo). DoG = s1 - s0 is formed in shared memory (never
// written to HBM), the squared gradient magnitude comes from the s0 tile.
// Phase A: every pixel takes the two CHEAP tests of the reference exactly: the magnitude test (:83) and the 25-tap sign count
// (:85-96, |pos - neg| <= 10), the latter from per-column counts of positive DoG values over the window's 5 rows (one small
// integer per tile position, shared by the 5 windows that contain the column). Both are exact, so there is no error bound to
// argue about; the sign count alone rejects ~80 % of all pixels (a window passes only where the zero crossing runs through it),
// the two together ~83 %. The survivors are compacted into a shared-memory list.
// Phase B: the threads work through the list densely and run the reference's own arithmetic (sequential dot products :98,
// tests :99-107), so what is accepted, and every accepted value, is bit-exact.
// (Round 1 filtered with a float estimate of the plane fit and a rigorous error bound: ~150 instructions per pixel in
// phase A because a warp runs the estimate when any of its lanes needs it, and three more shared planes to build. ncu, 64
// lanes x 752x480: 273 thread instructions per pixel, issue-bound at 82 %.)
// Accepted pixels set their bit in the per-row flag words and park (theta0, theta1, xs, ys) in the fit buffer for the emit
// pass. Also resets the id mask.
#define DT_W 64
#define DT_H 16
#define DT_PITCH (DT_W + 5)
__global__ void __launch_bounds__(256) k_detect_flags(const float* __restrict__ scales, const float* __restrict__ thr_dev,
                                                      unsigned* __restrict__ flags, int* __restrict__ rowcount, int* __restrict__ mask,
                                                      float4* __restrict__ fit, const __grid_constant__ DevParams P, int wpr) {
  __shared__ float s0t[DT_H + 4][DT_PITCH];
  __shared__ float dogt[DT_H + 4][DT_PITCH];
  __shared__ unsigned char colP[DT_H][DT_PITCH + 3];   // positives among the 5 window rows of a column (0..5)
  __shared__ unsigned short cand[DT_W * DT_H];
  __shared__ unsigned sflags[DT_H][2];
  __shared__ int ncand;
  const int b = blockIdx.z;
  const int rows = P.rows, cols = P.cols;
  const size_t N = (size_t)P.npix;
  const float* s0 = scales + (size_t)(b * 2) * N;
  const float* s1 = s0 + N;
  const int c0 = blockIdx.x * DT_W, r0 = blockIdx.y * DT_H;
  const int tid = threadIdx.x, lane = tid & 31;
  if ((cols & 3) == 0) {
    // aligned 16-byte loads covering columns [c0 - 4, c0 + 68): 18 float4 per tile row, all of a thread's loads in flight
    // before the first use (the tile load is the latency-critical part of this kernel)
    constexpr int V4 = (DT_W + 8) / 4, NV = (DT_H + 4) * V4;
    float4 a[2], d[2];
    int trj[2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int i = tid + k * 256;
      const int tr = i / V4, j = i % V4;
      const int r = r0 - 2 + tr, q4 = c0 - 4 + 4 * j;
      trj[k] = (i < NV) ? (tr << 8 | j) : -1;
      a[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      d[k] = a[k];
      if (i < NV && r >= 0 && r < rows && q4 >= 0 && q4 < cols) {
        a[k] = *reinterpret_cast<const float4*>(s0 + (size_t)r * cols + q4);
        d[k] = *reinterpret_cast<const float4*>(s1 + (size_t)r * cols + q4);
      }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      if (trj[k] < 0) continue;
      const int tr = trj[k] >> 8, tc0 = 4 * (trj[k] & 255) - 2;
      const float av[4] = {a[k].x, a[k].y, a[k].z, a[k].w};
      const float dv[4] = {d[k].x - a[k].x, d[k].y - a[k].y, d[k].z - a[k].z, d[k].w - a[k].w};   // calculateDoG, scale_space.cpp:216
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int tc = tc0 + e;
        if (tc >= 0 && tc < DT_W + 4) { s0t[tr][tc] = av[e]; dogt[tr][tc] = dv[e]; }
      }
    }
  } else {
    for (int i = tid; i < (DT_H + 4) * (DT_W + 4); i += 256) {
      const int tr = i / (DT_W + 4), tc = i % (DT_W + 4);
      const int r = r0 - 2 + tr, q = c0 - 2 + tc;
      float a = 0.0f, d = 0.0f;
      if (r >= 0 && r < rows && q >= 0 && q < cols) {
        a = s0[(size_t)r * cols + q];
        d = s1[(size_t)r * cols + q] - a;                                              // calculateDoG, scale_space.cpp:216
      }
      s0t[tr][tc] = a;
      dogt[tr][tc] = d;
    }
  }
  if (tid < DT_H * 2) sflags[tid >> 1][tid & 1] = 0u;
  if (tid == 0) ncand = 0;
  __syncthreads();
  {
    // colP[vr][vc] = positives among dogt[vr .. vr + 4][vc] (:92, NaN counts as negative). A thread takes four consecutive rows of one
    // column: eight loads and a sliding count instead of twenty loads (and no division of a flat index by the 68-column pitch)
    static_assert(DT_H == 16 && DT_W == 64, "colP mapping: 4 row groups x 64 columns per 256 threads, the 4 halo columns by 16 threads");
    auto col4 = [&](int vc, int g) {
      int pz[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) pz[k] = dogt[4 * g + k][vc] > 0.0f;
      int sct = pz[0] + pz[1] + pz[2] + pz[3] + pz[4];
      colP[4 * g][vc] = (unsigned char)sct;
#pragma unroll
      for (int k = 1; k < 4; ++k) { sct += pz[k + 4] - pz[k - 1]; colP[4 * g + k][vc] = (unsigned char)sct; }
    };
    col4(tid & 63, tid >> 6);
    if (tid < 16) col4(DT_W + (tid & 3), tid >> 2);
  }
  __syncthreads();
  const float thr = thr_dev[b];
  const float tm = thr * 765.0f;
  const float mag_threshold = tm * tm;                                                 // edge_detector.cpp:71
  const float tg = tm * P.dog_threshold;
  const float gts = tg * tg;                                                           // :70
  unsigned cm[(DT_W * DT_H) / 256];   // ballots of the thread's four pixels
#pragma unroll
  for (int i = 0; i < (DT_W * DT_H) / 256; ++i) {
    const int p = tid + i * 256;
    const int pr = p / DT_W, pc = p % DT_W;
    const int r = r0 + pr, q = c0 + pc;
    bool c = false;
    if (r < rows && q < cols) {
      mask[(size_t)b * N + (size_t)r * cols + q] = -1;
      if (r >= 2 && r < rows - 2 && q >= 2 && q < cols - 2) {
        const int tr = pr + 2, tc = pc + 2;
        const float dx = s0t[tr][tc + 1] - s0t[tr][tc - 1];
        const float dy = s0t[tr + 1][tc] - s0t[tr - 1][tc];
        const float mag = dx * dx + dy * dy;                                           // scale_space.cpp:228-230
        c = !(mag < mag_threshold);                                                    // :83
        if (c) {   // sign count of the 5x5 DoG window, edge_detector.cpp:85-96: pn = positives - negatives
          const int pos = (int)colP[pr][pc] + (int)colP[pr][pc + 1] + (int)colP[pr][pc + 2] + (int)colP[pr][pc + 3] + (int)colP[pr][pc + 4];
          c = !((float)abs(2 * pos - 25) > P.pn_threshold);                            // :96 (small integers: exact in float32 as in double)
        }
      }
    }
    cm[i] = __ballot_sync(0xFFFFFFFFu, c);
  }
  {   // one slot reservation per warp for its (up to 128) candidates; their order in the list does not matter (the fit below is per candidate)
    int tot = 0;
#pragma unroll
    for (int i = 0; i < (DT_W * DT_H) / 256; ++i) tot += __popc(cm[i]);
    int base = 0;
    if (lane == 0 && tot) base = atomicAdd(&ncand, tot);
    base = __shfl_sync(0xFFFFFFFFu, base, 0);
#pragma unroll
    for (int i = 0; i < (DT_W * DT_H) / 256; ++i) {
      if ((cm[i] >> lane) & 1u) cand[base + __popc(cm[i] & ((1u << lane) - 1u))] = (unsigned short)(tid + i * 256);
      base += __popc(cm[i]);
    }
  }
  __syncthreads();
  const int nc = ncand;
  for (int i = tid; i < nc; i += 256) {
    const int p = cand[i];
    const int pr = p / DT_W, pc = p % DT_W;
    float y[25];
#pragma unroll
    for (int dr = 0; dr < 5; ++dr)
#pragma unroll
      for (int dc = 0; dc < 5; ++dc) y[dr * 5 + dc] = dogt[pr + dr][pc + dc];
    FitResult fr;
    if (rbv_fit_plane<true>(y, P, gts, fr)) {
      atomicOr(&sflags[pr][pc >> 5], 1u << (pc & 31));
      fit[(size_t)b * N + (size_t)(r0 + pr) * cols + (c0 + pc)] = make_float4(fr.th0, fr.th1, fr.xs, fr.ys);
    }
  }
  __syncthreads();
  if (tid < DT_H * 2) {
    const int pr = tid >> 1, w = blockIdx.x * 2 + (tid & 1);
    const int r = r0 + pr;
    if (r < rows && w < wpr) flags[((size_t)b * rows + r) * wpr + w] = sflags[pr][tid & 1];
    if ((tid & 1) == 0 && r < rows) {
      const int cnt = __popc(sflags[pr][0]) + __popc(sflags[pr][1]);
      if (cnt) atomicAdd(&rowcount[b * rows + r], cnt);
    }
  }
}

// Pass 2: exclusive prefix of the per-row counts -> first keyline id of every row; count = min(total, keylines_max)
__global__ void __launch_bounds__(1024) k_detect_scan(int* __restrict__ rowcount, int* __restrict__ rowoffset, int* __restrict__ count,
                                                      int rows, int kmax) {
  __shared__ int part[1024];
  const int b = blockIdx.x, t = threadIdx.x;
  const int per = (rows + 1023) / 1024;
  const int lo = t * per, hi = min(lo + per, rows);
  int s = 0;
  for (int r = lo; r < hi; ++r) s += rowcount[b * rows + r];
  part[t] = s;
  __syncthreads();
  // Hillis-Steele inclusive scan over 1024 partials
  for (int off = 1; off < 1024; off <<= 1) {
    int v = (t >= off) ? part[t - off] : 0;
    __syncthreads();
    part[t] += v;
    __syncthreads();
  }
  int run = part[t] - s;
  for (int r = lo; r < hi; ++r) {
    rowoffset[b * rows + r] = run;
    run += rowcount[b * rows + r];
    rowcount[b * rows + r] = 0;   // consumed: left clean for the next frame's k_detect_flags (no fill launch per frame)
  }
  if (t == 1023) count[b] = min(part[1023], kmax);
}

// Pass 3: emit keylines in row-major id order (edge_detector.cpp:109-119, KeyLine ctor types/keyline.hpp:42-59).
// One warp per image row, EM_ROWS rows per block. Step 1: lane l owns flag word l (+32 per round) and lists the columns
// of its set bits in shared memory at (prefix popcount) — the row's keylines in id order. Step 2: lane j takes list entry
// j (+32 ...): consecutive lanes hold consecutive ids, so the plane-fit read-back (16 B per keyline, parked by pass 1)
// is issued for 32 keylines at once and every SoA store is coalesced.
#define EM_ROWS 4
#define EM_MAXW 1024   // columns per row the shared list can hold (wider images take the direct path below)
__global__ void __launch_bounds__(32 * EM_ROWS) k_detect_emit(const float4* __restrict__ fit, const unsigned* __restrict__ flags,
                                                              const int* __restrict__ rowoffset, KeylineSoA K, int* __restrict__ mask,
                                                              const __grid_constant__ DevParams P, int wpr, float4* __restrict__ tgt) {
  __shared__ unsigned short list[EM_ROWS][EM_MAXW];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int b = blockIdx.y, r = blockIdx.x * EM_ROWS + warp + 2;
  const int rows = P.rows, cols = P.cols, kmax = P.kmax;
  const size_t N = (size_t)P.npix;
  if (r >= rows - 2) return;
  const int base = rowoffset[b * rows + r];
  if (base >= P.keylines_max) return;
  int total = 0;
  for (int w0 = 0; w0 < wpr; w0 += 32) {
    const int w = w0 + lane;
    unsigned word = (w < wpr) ? flags[((size_t)b * rows + r) * wpr + w] : 0u;
    const int cnt = __popc(word);
    int inc = cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      int v = __shfl_up_sync(0xFFFFFFFFu, inc, off);
      if (lane >= off) inc += v;
    }
    int pos = total + inc - cnt;
    total += __shfl_sync(0xFFFFFFFFu, inc, 31);
    while (word) {
      const int bit = __ffs(word) - 1;
      word &= word - 1;
      list[warp][pos++] = (unsigned short)(w * 32 + bit);
    }
  }
  __syncwarp();
  total = min(total, P.keylines_max - base);                                            // :113-119 (keylines_max cut)
  for (int j = lane; j < total; j += 32) {
    const int q = list[warp][j];
    const int id = base + j;
    const unsigned pix = (unsigned)(r * cols + q);
    const float4 fr = fit[(size_t)b * N + pix];                                         // (theta0, theta1, xs, ys)
    const float px = (float)q + fr.z, py = (float)r + fr.w;                             // :109
    const size_t o = (size_t)b * kmax + id;
    K.pos[o] = make_float2(px, py);
    const float2 pi = make_float2(px - P.cx, py - P.cy);                                // :110
    K.pos_img[o] = pi;
    K.match_pos_img[o] = pi;
    K.grad[o] = make_float2(fr.x, fr.y);
    K.match_grad[o] = make_float2(0.0f, 0.0f);
    const float gn = sqrtf(fr.x * fr.x + fr.y * fr.y);
    K.gnorm[o] = gn;
    if (tgt) {   // the packed target record of k_pack_targets (rbv_tracker.cu), written here when the map's distance field follows its detect
      tgt[2 * o] = make_float4(fr.x, fr.y, px, py);
      tgt[2 * o + 1] = make_float4(fr.x / gn, fr.y / gn, gn, 0.0f);                       // core.cpp:64-65
    }
    K.match_gnorm[o] = 0.0f;
    K.rho[o] = RBV_RHO_INIT;
    K.srho[o] = RBV_RHO_MAX;
    K.id_prev[o] = -1; K.id_next[o] = -1; K.match_id[o] = -1; K.match_fwd[o] = -1; K.match_kf[o] = -1;
    K.matches[o] = 0u;
    K.pixel[o] = pix;
    K.residual[o] = 0.0f;
    mask[(size_t)b * N + pix] = id;                                                     // :111-112
  }
}

// the same for rows wider than the shared list: lane l walks the bits of its own flag words
__global__ void __launch_bounds__(32) k_detect_emit_wide(const float4* __restrict__ fit, const unsigned* __restrict__ flags,
                                                         const int* __restrict__ rowoffset, KeylineSoA K, int* __restrict__ mask,
                                                         const __grid_constant__ DevParams P, int wpr, float4* __restrict__ tgt) {
  const int b = blockIdx.y, r = blockIdx.x + 2, lane = threadIdx.x;
  const int rows = P.rows, cols = P.cols, kmax = P.kmax;
  const size_t N = (size_t)P.npix;
  int base = rowoffset[b * rows + r];
  if (base >= P.keylines_max) return;
  for (int w0 = 0; w0 < wpr; w0 += 32) {
    const int w = w0 + lane;
    unsigned word = (w < wpr) ? flags[((size_t)b * rows + r) * wpr + w] : 0u;
    const int cnt = __popc(word);
    int inc = cnt;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      int v = __shfl_up_sync(0xFFFFFFFFu, inc, off);
      if (lane >= off) inc += v;
    }
    int id = base + inc - cnt;
    base += __shfl_sync(0xFFFFFFFFu, inc, 31);
    while (word) {
      const int bit = __ffs(word) - 1;
      word &= word - 1;
      if (id < P.keylines_max) {
        const int q = w * 32 + bit;
        const unsigned pix = (unsigned)(r * cols + q);
        const float4 fr = fit[(size_t)b * N + pix];
        const float px = (float)q + fr.z, py = (float)r + fr.w;
        const size_t o = (size_t)b * kmax + id;
        K.pos[o] = make_float2(px, py);
        const float2 pi = make_float2(px - P.cx, py - P.cy);
        K.pos_img[o] = pi;
        K.match_pos_img[o] = pi;
        K.grad[o] = make_float2(fr.x, fr.y);
        K.match_grad[o] = make_float2(0.0f, 0.0f);
        const float gn = sqrtf(fr.x * fr.x + fr.y * fr.y);
        K.gnorm[o] = gn;
        if (tgt) {   // packed target record, as in k_detect_emit
          tgt[2 * o] = make_float4(fr.x, fr.y, px, py);
          tgt[2 * o + 1] = make_float4(fr.x / gn, fr.y / gn, gn, 0.0f);
        }
        K.match_gnorm[o] = 0.0f;
        K.rho[o] = RBV_RHO_INIT;
        K.srho[o] = RBV_RHO_MAX;
        K.id_prev[o] = -1; K.id_next[o] = -1; K.match_id[o] = -1; K.match_fwd[o] = -1; K.match_kf[o] = -1;
        K.matches[o] = 0u;
        K.pixel[o] = pix;
        K.residual[o] = 0.0f;
        mask[(size_t)b * N + pix] = id;
      }
      ++id;
    }
  }
}

// joinEdges + nextKeylineIdx (edge_detector.cpp:125-165). id_prev: the sequential loop lets the largest idx win -> atomicMax.
__global__ void k_join(KeylineSoA K, const int* __restrict__ count, const int* __restrict__ mask, const __grid_constant__ DevParams P) {
  const int b = blockIdx.y, idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= count[b]) return;
  const size_t o = (size_t)b * P.kmax + idx;
  const float2 p = K.pos[o], g = K.grad[o];
  const int x = rbv_d2i((double)p.x + 0.5), y = rbv_d2i((double)p.y + 0.5);             // :128-129
  const float tx = -g.y, ty = g.x;
  const int sx = (ty > 0.0f) ? ((tx > 0.0f) ? 1 : -1) : ((tx < 0.0f) ? -1 : 1);
  const int sy = (ty > 0.0f) ? 1 : -1;
  const int* m = mask + (size_t)b * P.npix;
  int hit = m[y * P.cols + x + sx];
  if (hit < 0) hit = m[(y + sy) * P.cols + x];
  if (hit < 0) hit = m[(y + sy) * P.cols + x + sx];
  if (hit < 0) return;
  atomicMax(&K.id_prev[(size_t)b * P.kmax + hit], idx);                                // :133
  K.id_next[o] = hit;                                                                  // :134
}

// tuneThreshold (edge_detector.cpp:167-186). The histogram scan at :183 always runs to i == num_bins (bin 0 holds the
// maximum, so bins 1.. can never reach keylines_max), hence threshold = max - float(100*(max-min))/100.
__global__ void __launch_bounds__(1024) k_tune(KeylineSoA K, const int* __restrict__ count, float* __restrict__ threshold, int kmax) {
  __shared__ float smx[32], smn[32];
  const int b = blockIdx.x, n = count[b], t = threadIdx.x;
  float mx = -INFINITY, mn = INFINITY;
  for (int i = t; i < n; i += 1024) {
    const float g = K.gnorm[(size_t)b * kmax + i];
    mx = fmaxf(mx, g);
    mn = fminf(mn, g);
  }
#pragma unroll
  for (int off = 16; off; off >>= 1) {
    mx = fmaxf(mx, __shfl_xor_sync(0xFFFFFFFFu, mx, off));
    mn = fminf(mn, __shfl_xor_sync(0xFFFFFFFFu, mn, off));
  }
  if ((t & 31) == 0) { smx[t >> 5] = mx; smn[t >> 5] = mn; }
  __syncthreads();
  if (t == 0) {
    for (int i = 1; i < 32; ++i) { mx = fmaxf(mx, smx[i]); mn = fminf(mn, smn[i]); }
    if (n < 1) {
      threshold[b] = -1.0f;  // reference: undefined behaviour on an empty map; the map keeps its initial -1
    } else {
      const float span = (float)RBV_TUNE_BINS * (mx - mn);                             // i*(max_dog-min_dog), i == num_bins
      threshold[b] = mx - span / (float)RBV_TUNE_BINS;                                 // :184
    }
  }
}

// EdgeDetector::buildEdgeMap + joinEdges + tuneThreshold on the scale space already in bufA
int fe_detect(rbv_ctx* c, rbv_map* map, float4* tgt) {
  const DevParams& P = c->P;
  const int B = P.lanes, rows = P.rows, cols = P.cols;
  const int wpr = (cols + 31) / 32;
  // c->rowcount is all zero here: cleared at creation and by every k_detect_scan after it has consumed the counts
  dim3 g1((cols + DT_W - 1) / DT_W, (rows + DT_H - 1) / DT_H, B);
  float4* fit = reinterpret_cast<float4*>(c->rowsum);   // rowsum + bufB are one contiguous 4*B*N float scratch, free after the scale space
  RBV_LAUNCH(c, k_detect_flags, (g1), (256), 0, c->bufA, c->thr_dev, c->flags, c->rowcount, map->mask, fit, P, wpr);
  RBV_LAUNCH(c, k_detect_scan, (B), (1024), 0, c->rowcount, c->rowoffset, map->count, rows, P.keylines_max);
  if (cols <= EM_MAXW)
    RBV_LAUNCH(c, k_detect_emit, (dim3((rows - 4 + EM_ROWS - 1) / EM_ROWS, B)), (32 * EM_ROWS), 0, fit, c->flags, c->rowoffset, map->k, map->mask, P, wpr, tgt);
  else
    RBV_LAUNCH(c, k_detect_emit_wide, (dim3(rows - 4, B)), (32), 0, fit, c->flags, c->rowoffset, map->k, map->mask, P, wpr, tgt);
  RBV_LAUNCH(c, k_join, (dim3((P.kmax + 255) / 256, B)), (256), 0, map->k, map->count, map->mask, P);
  RBV_LAUNCH(c, k_tune, (B), (1024), 0, map->k, map->count, map->threshold, P.kmax);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// debug taps: DoG (which = 2) / squared gradient magnitude (which = 3) of one lane, recomputed from the scales
__global__ void k_dog_mag(const float* __restrict__ s0, const float* __restrict__ s1, float* __restrict__ out, int rows, int cols, int which) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (q >= cols) return;
  const size_t o = (size_t)r * cols + q;
  if (which == 2) { out[o] = s1[o] - s0[o]; return; }
  float v = 0.0f;
  if (r >= 1 && r < rows - 1 && q >= 1 && q < cols - 1) {
    const float dx = s0[o + 1] - s0[o - 1], dy = s0[o + cols] - s0[o - cols];
    v = dx * dx + dy * dy;
  }
  out[o] = v;
}
int fe_dog_mag_debug(rbv_ctx* c, int lane, int which, float* d_out) {
  const size_t N = (size_t)c->P.npix;
  const float* s0 = c->bufA + (size_t)(lane * 2) * N;
  RBV_LAUNCH(c, k_dog_mag, (dim3((c->P.cols + 127) / 128, c->P.rows)), (128), 0, s0, s0 + N, d_out, c->P.rows, c->P.cols, which);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// rbv_box_pass_noalloc: one box pass (createIntegralImage + average) of a device image without any allocation.
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code under
// CUDA-graph capture, tight serving loops, harnesses that forbid it). It is the split path's k_rowscan(4) + k_colbox pair - the same
// kernels, unchanged, that rbv_debug_box_pass runs by default - with the row-prefix image, the scratch between the two, in a
// file-scope device array instead of the context's cudaMalloc'd `rowsum`. Every device that loads the library holds the array
// (RBV_NOALLOC_MAX_PIXELS floats = 4 MiB) until exit.
// The array's NAME is a host-side stand-in in host code, not a device address: the address is taken once per device, with that
// device current, by cudaGetSymbolAddress, and that pointer is what the kernels get.
// All calls share the array and it keeps the last call's contents; nothing has to start at zero (k_rowscan writes every word of
// rows x cols before k_colbox reads it). Two calls that overlap on the GPU (different streams, no ordering between them) race on it.
static __device__ __align__(256) float rbv_noalloc_rowsum[RBV_NOALLOC_MAX_PIXELS];

static std::mutex noalloc_mu;
static float* noalloc_ptr[64];   // device address of rbv_noalloc_rowsum, per device ordinal

static int noalloc_scratch(float** out) {
  int dev = -1;
  RBV_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { rbv_set_error("rbv_box_pass_noalloc: device ordinal %d out of range", dev); return RBV_ERR_ARG; }
  std::lock_guard<std::mutex> lk(noalloc_mu);
  if (!noalloc_ptr[dev]) {   // set-up, once per device, that device current
    float* p = nullptr;
    RBV_CUDA(cudaGetSymbolAddress((void**)&p, rbv_noalloc_rowsum));
    if (!p) { rbv_set_error("rbv_box_pass_noalloc: cudaGetSymbolAddress returned a null address"); return RBV_ERR_CUDA; }
    noalloc_ptr[dev] = p;
  }
  *out = noalloc_ptr[dev];
  return RBV_OK;
}

extern "C" int rbv_box_pass_noalloc(const float* d_in, float* d_out, int rows, int cols, int width, void* cuda_stream) {
  if (!d_in || !d_out) { rbv_set_error("rbv_box_pass_noalloc: null device pointer"); return RBV_ERR_ARG; }
  if (rows < 1 || cols < 1 || rows > 32767 || cols > 32767) { rbv_set_error("rbv_box_pass_noalloc: bad image size %d x %d", rows, cols); return RBV_ERR_ARG; }
  if ((long long)rows * cols > (long long)RBV_NOALLOC_MAX_PIXELS) {
    rbv_set_error("rbv_box_pass_noalloc: %d x %d = %lld pixels exceed the static scratch of RBV_NOALLOC_MAX_PIXELS = %d pixels", rows, cols,
                  (long long)rows * cols, (int)RBV_NOALLOC_MAX_PIXELS);
    return RBV_ERR_ARG;
  }
  if (width < 3 || width > 15 || (width & 1) == 0) { rbv_set_error("rbv_box_pass_noalloc: box width %d not an odd number in 3..15", width); return RBV_ERR_ARG; }
  float* scratch = nullptr;
  { const int rc = noalloc_scratch(&scratch); if (rc != RBV_OK) return rc; }
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const size_t N = (size_t)rows * cols;
  if ((cols & 3) == 0 && ((uintptr_t)d_in & 15) == 0)
    k_rowscan4<<<dim3((rows + 31) / 32, 1), 32, 0, st>>>(d_in, scratch, rows, cols, N, N);
  else
    k_rowscan<<<dim3((rows + 31) / 32, 1), 32, 0, st>>>(d_in, scratch, rows, cols, N, N);
  const int W = CB_TPB - (2 * (width / 2) + 1);
  k_colbox<<<dim3((cols + W - 1) / W, 1), CB_TPB, 0, st>>>(scratch, d_out, d_out, rows, cols, N, N, 1, width, width, 0);
  RBV_CUDA(cudaGetLastError());
  return RBV_OK;
}
