// tridiag_cr_tma.cu -- batched cyclic reduction, level-major operands, one matrix and one right-hand side per
// column, many columns (BASELINE config 3: cr over 345,600 columns x 128 levels).
//
// Replaces the per-column loops around ekat::tridiag::cr(team, dl, d, du, X) (reference
// src/algorithm/ekat_tridiag.hpp:385-458). Update formulas, operand order and the clamped-neighbour trick
// (:408-418) are kept literally and unfused, every * + / separately rounded, so the results are bit-identical to
// the reference's CR at -ffp-contract=off whatever the thread geometry (rows touched in a level are never read in
// that level).
//
// A CTA owns a tile of 16 columns: the four operand planes arrive by one 2D TMA tensor copy each as
// [level][16 columns] (64 KB for 128 levels f64, so three CTAs share an SM and one tile's loads and stores overlap
// the others' arithmetic); a thread is (row slot, column): lane%16 = column, the 16 row slots of the 8 warps take
// the rows of a level round-robin exactly as the reference's team does (`for (i = stride*team_id; i < nrow;
// i += stride*nteam)`, :403). The reference's team_barrier per level (:422, :456) is a __syncthreads. The solution
// plane leaves by a TMA tensor store. HBM traffic is the algorithmic minimum (read dl, d, du, X once, write X once).
#include <cuda.h>

#include "tridiag_tma.cuh"

namespace ekb {

constexpr int kCrWarps = 8, kCrTW = 16, kCrSlots = kCrWarps * 32 / kCrTW;

template <typename R>
__global__ void __launch_bounds__(32 * kCrWarps)
k_tridiag_cr_tma(const __grid_constant__ CUtensorMap mdl, const __grid_constant__ CUtensorMap md,
                 const __grid_constant__ CUtensorMap mdu, const __grid_constant__ CUtensorMap mx, int nrow) {
  constexpr int TW = kCrTW;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem_raw);
  const size_t plane = (((size_t)nrow * TW * sizeof(R) + 127) & ~(size_t)127) / sizeof(R);  // TMA boxes land 128-byte aligned
  R* sdl = reinterpret_cast<R*>(smem_raw + 128);
  R* sd = sdl + plane;
  R* sdu = sd + plane;
  R* sx = sdu + plane;

  const int tid = threadIdx.x;
  const int c = tid % TW, w = tid / TW;  // column of the tile, row slot
  const int c0 = blockIdx.x * TW;
  if (tid == 0) {
    mbar_init(bar, 1);
    fence_mbar_init();
    mbar_expect_tx(bar, (unsigned)(4 * (size_t)nrow * TW * sizeof(R)));  // one box of nrow (<= 256) levels per plane
    tma_load_2d(sdl, &mdl, c0, 0, bar);
    tma_load_2d(sd, &md, c0, 0, bar);
    tma_load_2d(sdu, &mdu, c0, 0, bar);
    tma_load_2d(sx, &mx, c0, 0, bar);
  }
  __syncthreads();
  mbar_wait(bar, 0u);

  R* pl = sdl + c;
  R* pd = sd + c;
  R* pu = sdu + c;
  R* px = sx + c;
  constexpr int L = TW;  // stride between levels of one column
  int os = 1, stride;
  while ((stride = (os << 1)) < nrow) {  // down-sweep :401-423
    for (int i = stride * w; i < nrow; i += stride * kCrSlots) {
      const bool im_ok = i - os >= 0, ip_ok = i + os < nrow;
      const int im = im_ok ? i - os : i, ip = ip_ok ? i + os : i;  // :413-414
      const R f1 = im_ok ? x_div(-pl[i * L], pd[im * L]) : (R)0;   // :408-409
      const R f2 = ip_ok ? x_div(-pu[i * L], pd[ip * L]) : (R)0;
      px[i * L] = x_add(px[i * L], x_add(x_mul(f1, px[im * L]), x_mul(f2, px[ip * L])));  // :418
      pl[i * L] = x_mul(f1, pl[im * L]);                                                   // :415-417, statement order kept
      pu[i * L] = x_mul(f2, pu[ip * L]);
      pd[i * L] = x_add(pd[i * L], x_add(x_mul(f1, pu[im * L]), x_mul(f2, pl[ip * L])));
    }
    os <<= 1;
    __syncthreads();  // team_barrier :422
  }
  if (w == 0) {  // bottom 1x1 or 2x2 solve :427-437
    if (os >= nrow) {
      px[0] = x_div(px[0], pd[0]);
    } else {
      const R d0 = pd[0], dos = pd[os * L], du0 = pu[0], dlos = pl[os * L];
      const R det = x_sub(x_mul(d0, dos), x_mul(du0, dlos));
      const R x0 = px[0], x1 = px[os * L];
      px[0] = x_div(x_sub(x_mul(dos, x0), x_mul(du0, x1)), det);
      px[os * L] = x_div(x_sub(x_mul(d0, x1), x_mul(dlos, x0)), det);
    }
  }
  __syncthreads();  // :438
  os >>= 1;
  while (os) {  // up-sweep :442-457
    stride = os << 1;
    for (int i = stride * w + os; i < nrow; i += stride * kCrSlots) {
      const int im = i - os, ip = i + os;
      const bool im_ok = im >= 0, ip_ok = ip < nrow;
      R f = 0;
      f = x_add(f, im_ok ? x_mul(pl[i * L], px[im * L]) : (R)0);  // :450-451
      f = x_add(f, ip_ok ? x_mul(pu[i * L], px[ip * L]) : (R)0);
      px[i * L] = x_div(x_sub(px[i * L], f), pd[i * L]);          // :452
    }
    os >>= 1;
    fence_proxy_async();  // (the last level's writes are what the tensor store reads)
    __syncthreads();      // :456
  }
  if (nrow <= 2) {  // no up-sweep level ran: the bottom solve's writes still need the proxy fence
    fence_proxy_async();
    __syncthreads();
  }
  if (tid == 0) {
    tma_store_2d(&mx, c0, 0, sx);
    tma_store_commit();
    tma_store_wait_read<0>();
  }
}

// Returns EKAT_B200_ERR_UNSUPPORTED when the operands do not meet TMA's requirements or a tile does not fit on chip;
// the caller then uses the generic tile kernel (tridiag_tile.cu).
template <typename R>
int tridiag_cr_tma(R* dl, R* d, R* du, R* X, int64_t ncol, int nlev, int64_t a_stride, int64_t x_stride, cudaStream_t stream) {
  auto al = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  if (!al(dl) || !al(d) || !al(du) || !al(X) || (a_stride * sizeof(R)) % 16 || (x_stride * sizeof(R)) % 16)
    return EKAT_B200_ERR_UNSUPPORTED;
  if (ncol >= (int64_t)1 << 31 || ncol < 4 * kNumSms * kCrTW) return EKAT_B200_ERR_UNSUPPORTED;  // few columns: rows over many threads instead
  if (nlev > 256) return EKAT_B200_ERR_UNSUPPORTED;  // one TMA box per plane
  const size_t smem = 128 + 4 * (((size_t)nlev * kCrTW * sizeof(R) + 127) & ~(size_t)127);
  if (smem > kMaxSmemPerCta / 2) return EKAT_B200_ERR_UNSUPPORTED;  // (long columns: the tile kernel gives a column more threads)
  CUtensorMap mdl, md, mdu, mx;
  const int boxr = nlev;
  if (!make_tensor_map_2d<R>(&mdl, dl, ncol, nlev, a_stride, kCrTW, boxr, 0) || !make_tensor_map_2d<R>(&md, d, ncol, nlev, a_stride, kCrTW, boxr, 0) ||
      !make_tensor_map_2d<R>(&mdu, du, ncol, nlev, a_stride, kCrTW, boxr, 0) || !make_tensor_map_2d<R>(&mx, X, ncol, nlev, x_stride, kCrTW, boxr, 0))
    return EKAT_B200_ERR_UNSUPPORTED;
  auto k = k_tridiag_cr_tma<R>;
  int rc = allow_smem(k, smem);
  if (rc) return rc;
  k<<<(unsigned)((ncol + kCrTW - 1) / kCrTW), 32 * kCrWarps, smem, stream>>>(mdl, md, mdu, mx, nlev);
  note_launch(sizeof(R) == 8 ? "tridiag_cr_tma<f64>" : "tridiag_cr_tma<f32>");
  return check_cuda(cudaGetLastError());
}

template int tridiag_cr_tma<double>(double*, double*, double*, double*, int64_t, int, int64_t, int64_t, cudaStream_t);
template int tridiag_cr_tma<float>(float*, float*, float*, float*, int64_t, int, int64_t, int64_t, cudaStream_t);

}  // namespace ekb
