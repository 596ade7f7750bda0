// Resident whole-loop kernel: ONE launch per forward_simulate / adjoint_simulate call, for grids small enough that a
// time step is shorter than the gap between two kernel launches (the reference's default 250 x 150 configuration:
// loops src/fdModel.cpp:552-751 and :814-958).
//
// The grid is cut into strips of `rows` consecutive rows, one CTA per strip, all CTAs co-resident (cooperative
// launch).  A CTA runs the WHOLE time loop for its strip.  Per time step it
//   1. waits until the CTAs that own rows within 4 of its strip have published the previous step (one 32-bit
//      counter per CTA in global memory, st.release / ld.acquire at gpu scope: neighbours only, no grid-wide barrier),
//   2. pulls the old state it needs -- vx vz rows [xs-4, xe+4), txx tzz txz rows [xs-2, xe+2), each ONE contiguous
//      block because rows are contiguous in z -- from L2 into shared memory with five bulk-asynchronous copies
//      (cp.async.bulk + mbarrier),
//   3. stores the snapshot (:555-572) / correlates with it (:817-861) / samples the receivers (:576-596) on its own
//      rows from those tiles,
//   4. updates the stresses of rows [xs-2, xe+2) in place in shared memory (the two halo rows on each side are
//      recomputed, like the strips of fused.cuh) and, after a block barrier, the velocities of its own rows; applies
//      moment-tensor / adjoint-source injections (:681-751, :951-958) to the new velocities in list order,
//   5. writes its rows of the new state to the other ping-pong buffer in global memory and bumps its counter.
// lm, mu and 1/rho of the strip stay in shared memory for the whole loop.  The arithmetic is the shared point-wise
// code of stencil_math.cuh in the same order as every other variant, so "exact" results are bit-identical to the
// reference and the default arithmetic matches the other variants bit for bit.
//
// Why this is safe without a grid barrier: state S_t lives in buffer t&1.  A CTA starts step t only after its
// neighbours published S_t, i.e. finished step t-1 including all their reads of S_{t-1} from buffer (t+1)&1 -- the
// buffer the CTA overwrites with S_{t+1} at the end of step t.
#pragma once

#include "fused.cuh"

namespace psv {

constexpr unsigned kResidentSpinLimit = 1u << 22;

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(unsigned *p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(unsigned bar, unsigned parity) {
  unsigned ok;
  asm volatile(
      "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}

template <class M>
__global__ void __launch_bounds__(kResidentThreads, 1) resident_loop_kernel(const ResidentArgs a) {
  extern __shared__ __align__(128) unsigned char psv_smem[];
  double *sm = reinterpret_cast<double *>(psv_smem);
  __shared__ int s_abort;
  const FusedCommon &c = a.c;
  const FusedModel &m = a.models[blockIdx.y];
  const int tid = threadIdx.x, nthreads = blockDim.x;
  const int ld = c.g.ld, P2 = ld >> 1, nz = c.g.nz, nx = c.g.nx;
  const int gx0 = c.g.x_off, gnx = c.g.nx_global;
  const int R = a.rows;
  const int xs = blockIdx.x * R, xe = min(xs + R, nx);
  const ResidentLayout L(ld, R);
  double2 *tvx = reinterpret_cast<double2 *>(sm + L.vx), *tvz = reinterpret_cast<double2 *>(sm + L.vz);
  double2 *txx = reinterpret_cast<double2 *>(sm + L.xx), *tzz = reinterpret_cast<double2 *>(sm + L.zz);
  double2 *txz = reinterpret_cast<double2 *>(sm + L.xz);
  double2 *tlm = reinterpret_cast<double2 *>(sm + L.lm), *tmu = reinterpret_cast<double2 *>(sm + L.mu);
  double2 *tb = reinterpret_cast<double2 *>(sm + L.b);
  const unsigned bar = smem_u32(sm);
  // tile row of grid row j: velocities (j - xs + 4), stresses / lm / mu (j - xs + 2), b (j - xs)
#define VROW(j) ((j) - xs + 4)
#define SROW(j) ((j) - xs + 2)

  if (tid == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    s_abort = 0;
  }
  // the medium of the strip: resident for the whole loop
  {
    const double2 *glm = reinterpret_cast<const double2 *>(m.lm) + (long long)(xs - 2) * P2;
    const double2 *gmu = reinterpret_cast<const double2 *>(m.mu) + (long long)(xs - 2) * P2;
    const double2 *gb = reinterpret_cast<const double2 *>(m.b) + (long long)xs * P2;
    for (int i = tid; i < (xe - xs + 4) * P2; i += nthreads) {
      tlm[i] = __ldg(glm + i);
      tmu[i] = __ldg(gmu + i);
    }
    for (int i = tid; i < (xe - xs) * P2; i += nthreads) tb[i] = __ldg(gb + i);
  }
  // point operations of this strip (row-sorted lists: a contiguous range each)
  const bool fwd = a.direction > 0;
  int rq0 = 0, rq1 = 0, oq0 = 0, oq1 = 0;
  if (c.point_mode != POINT_NONE) {
    rq0 = c.rec_row_start[xs];
    rq1 = c.rec_row_start[xe];
  }
  if (c.point_mode == POINT_FORWARD) {
    oq0 = c.op_row_start[xs];
    oq1 = c.op_row_start[xe];
  }
  const bool inject = fwd ? oq1 > oq0 : (c.point_mode == POINT_ADJOINT && rq1 > rq0);
  // neighbours whose rows this strip reads: owners of rows [xs-4, xe+4) other than this CTA
  const int nb_lo = max(xs - 4, 0) / R, nb_hi = min(xe + 3, nx - 1) / R;
  const int vrows = xe - xs + 8, srows = xe - xs + 4;
  const unsigned vbytes = (unsigned)vrows * ld * 8u, sbytes = (unsigned)srows * ld * 8u;
  // this thread's injection operation (forward) / receiver (adjoint): constant over the loop
  RowOp op{};
  int my_rank = -1, my_x = 0, my_z = 0;
  size_t my_trace = 0;
  if (fwd) {
    if (oq0 + tid < oq1) {
      op = c.ops[oq0 + tid];
      my_rank = a.op_rank[oq0 + tid];
      my_trace = (size_t)op.source * c.nt;
    }
  } else if (c.point_mode == POINT_ADJOINT && rq0 + tid < rq1) {
    my_rank = a.rec_rank[rq0 + tid];
    my_x = c.rec_x[rq0 + tid];
    my_z = c.rec_z[rq0 + tid];
    my_trace = (size_t)c.rec_id[rq0 + tid] * c.nt;
  }
  __syncthreads();

  for (int step = 0; step < a.nsteps; ++step) {
    const int it = fwd ? step : a.nsteps - 1 - step;
    const int p = step & 1;
    const bool snap_step = it % a.si == 0;
    // this step's injection amplitudes: independent of the neighbours, fetched while waiting for them
    double inj_a = 0.0, inj_b = 0.0;
    if (my_rank >= 0) {
      if (fwd) {
        inj_a = c.stf[my_trace + it];
      } else {
        inj_a = m.ax[my_trace + it];
        inj_b = m.az[my_trace + it];
      }
    }
    // 1. wait for the neighbours' previous step
    if (step > 0) {
      const int nb = nb_lo + tid;
      if (nb <= nb_hi && nb != (int)blockIdx.x) {
        const unsigned *f = a.flags + 32 * (size_t)(blockIdx.y * gridDim.x + nb);
        unsigned spins = 0;
        while (ld_acquire_u32(f) < (unsigned)step) {
          if (++spins > kResidentSpinLimit || *reinterpret_cast<volatile unsigned *>(a.abort)) {
            *reinterpret_cast<volatile unsigned *>(a.abort) = 1u;
            s_abort = 1;
            break;
          }
        }
      }
    }
    __syncthreads();
    if (s_abort) return;
    // 2. old state of the strip + halos: five contiguous blocks, L2 -> shared memory
    if (tid == 0) {
      asm volatile("fence.proxy.async;" ::: "memory");  // generic-proxy writes (tiles, neighbours' rows) before the bulk copies
      mbar_expect_tx(bar, 2u * vbytes + 3u * sbytes);
      const long long voff = (long long)(xs - 4) * ld, soff = (long long)(xs - 2) * ld;
      bulk_g2s(smem_u32(tvx), m.f[p][0] + voff, vbytes, bar);
      bulk_g2s(smem_u32(tvz), m.f[p][1] + voff, vbytes, bar);
      bulk_g2s(smem_u32(txx), m.f[p][2] + soff, sbytes, bar);
      bulk_g2s(smem_u32(tzz), m.f[p][3] + soff, sbytes, bar);
      bulk_g2s(smem_u32(txz), m.f[p][4] + soff, sbytes, bar);
    }
    {
      unsigned spins = 0;
      while (!mbar_try_wait(bar, (unsigned)(step & 1))) {
        if (++spins > kResidentSpinLimit) {
          *reinterpret_cast<volatile unsigned *>(a.abort) = 1u;
          break;
        }
      }
    }
    // 3. side effects on the state before the update, own rows
    if (snap_step && (fwd ? a.store != 0 : true)) {
      const size_t snap_off = (size_t)(it / a.si) * a.grid_elems;
      if (fwd) {
        double2 *s = reinterpret_cast<double2 *>(m.snap + snap_off);
        const size_t fs2 = c.snap_field_stride >> 1;
        for (int i = tid; i < (xe - xs) * P2; i += nthreads) {
          const size_t g = (size_t)xs * P2 + i;
          s[g] = tvx[4 * P2 + i];
          s[g + fs2] = tvz[4 * P2 + i];
          s[g + 2 * fs2] = txx[2 * P2 + i];
          s[g + 3 * fs2] = tzz[2 * P2 + i];
          s[g + 4 * fs2] = txz[2 * P2 + i];
        }
      } else {
        const int cx0 = max(xs, c.wx0), cx1 = min(xe, c.wx1);
        const int kp0 = c.wz0 >> 1, kp1 = (c.wz1 + 1) >> 1;  // pairs that hold a window column
        const int npair = kp1 - kp0;
        const double2 *s = reinterpret_cast<const double2 *>(m.snap + snap_off);
        const size_t fs2 = c.snap_field_stride >> 1;
        for (int i = tid; i < (cx1 - cx0) * npair; i += nthreads) {
          const int j = cx0 + i / npair, kp = kp0 + i % npair, z0 = 2 * kp;
          const size_t g = (size_t)j * P2 + kp;
          const double2 fvx = __ldg(s + g), fvz = __ldg(s + g + fs2), fxx = __ldg(s + g + 2 * fs2),
                        fzz = __ldg(s + g + 3 * fs2), fxz = __ldg(s + g + 4 * fs2);
          double2 kr = *(reinterpret_cast<double2 *>(m.k_rho) + g), kl = *(reinterpret_cast<double2 *>(m.k_lam) + g),
                  km = *(reinterpret_cast<double2 *>(m.k_mu) + g);
          const double2 vx0 = tvx[VROW(j) * P2 + kp], vz0 = tvz[VROW(j) * P2 + kp];
          const double2 oxx = txx[SROW(j) * P2 + kp], ozz = tzz[SROW(j) * P2 + kp], oxz = txz[SROW(j) * P2 + kp];
          const double2 lm = tlm[SROW(j) * P2 + kp], mu = tmu[SROW(j) * P2 + kp];
          if (z0 >= c.wz0 && z0 < c.wz1)
            kernel_accumulate<M>(c.sidt, lm.x, mu.x, fvx.x, fvz.x, fxx.x, fzz.x, fxz.x, vx0.x, vz0.x, oxx.x, ozz.x,
                                 oxz.x, kr.x, kl.x, km.x);
          if (z0 + 1 >= c.wz0 && z0 + 1 < c.wz1)
            kernel_accumulate<M>(c.sidt, lm.y, mu.y, fvx.y, fvz.y, fxx.y, fzz.y, fxz.y, vx0.y, vz0.y, oxx.y, ozz.y,
                                 oxz.y, kr.y, kl.y, km.y);
          *(reinterpret_cast<double2 *>(m.k_rho) + g) = kr;
          *(reinterpret_cast<double2 *>(m.k_lam) + g) = kl;
          *(reinterpret_cast<double2 *>(m.k_mu) + g) = km;
        }
      }
      __syncthreads();  // the in-place stress update below overwrites what these threads read
    }
    if (fwd && c.point_mode == POINT_FORWARD) {  // receiver increments from the old velocities (prefix sum after the loop)
      for (int q = rq0 + tid; q < rq1; q += nthreads) {
        const int j = c.rec_x[q], z = c.rec_z[q];
        const double *rvx = reinterpret_cast<const double *>(tvx + VROW(j) * P2), *rvz = reinterpret_cast<const double *>(tvz + VROW(j) * P2);
        const double pvx = rvx[z], pvz = rvz[z];
        const size_t t = (size_t)c.rec_id[q] * c.nt + it;
        m.ux[t] = M::kExact ? __ddiv_rn(__dmul_rn(c.dt, pvx), c.dxdz) : (c.dt * pvx) * c.rdxdz;
        m.uz[t] = M::kExact ? __ddiv_rn(__dmul_rn(c.dt, pvz), c.dxdz) : (c.dt * pvz) * c.rdxdz;
      }
    }
    // 4a. stresses of rows [xs-2, xe+2), in place
    double2 *nf2 = reinterpret_cast<double2 *>(m.f[p ^ 1][2]), *nf3 = reinterpret_cast<double2 *>(m.f[p ^ 1][3]),
            *nf4 = reinterpret_cast<double2 *>(m.f[p ^ 1][4]);
    for (int i = tid; i < srows * P2; i += nthreads) {
      const int jr = i / P2, kp = i - jr * P2, j = xs - 2 + jr, z0 = 2 * kp;
      const bool rowok = gx0 + j >= 2 && gx0 + j < gnx - 2;
      const bool za = z0 >= 2 && z0 < nz - 2, zb = z0 + 1 >= 2 && z0 + 1 < nz - 2;
      if (!rowok || !(za || zb)) continue;  // frame rows / pad pairs keep their values
      const int v = (jr + 2) * P2 + kp;     // velocity-tile index of (j, kp)
      const double2 vx_m1 = tvx[v - P2], vx_0 = tvx[v], vx_p1 = tvx[v + P2], vx_p2 = tvx[v + 2 * P2];
      const double2 vxL = tvx[v - 1], vxR = tvx[v + 1];
      const double2 vz_m2 = tvz[v - 2 * P2], vz_m1 = tvz[v - P2], vz_0 = tvz[v], vz_p1 = tvz[v + P2];
      const double2 vzL = tvz[v - 1], vzR = tvz[v + 1];
      const double2 lm = tlm[i], mu = tmu[i];
      double2 sxx = txx[i], szz = tzz[i], sxz = txz[i];
      const double tr = taper_axis(c.taper_table, c.g.np, gx0 + j, gnx);
      if (za) {
        const double dxp_vx = diff4<M>(c.k, vx_p1.x, vx_0.x, vx_m1.x, vx_p2.x);
        const double dzm_vz = diff4<M>(c.k, vz_0.x, vzL.y, vzL.x, vz_0.y);
        const double dzp_vx = diff4<M>(c.k, vx_0.y, vx_0.x, vxL.y, vxR.x);
        const double dxm_vz = diff4<M>(c.k, vz_0.x, vz_m1.x, vz_m2.x, vz_p1.x);
        stress_update<M>(c.k, dxp_vx, dzm_vz, dzp_vx, dxm_vz, lm.x, mu.x,
                         dmin2(tr, taper_axis(c.taper_table, c.g.np, z0, nz)), sxx.x, szz.x, sxz.x);
      }
      if (zb) {
        const double dxp_vx = diff4<M>(c.k, vx_p1.y, vx_0.y, vx_m1.y, vx_p2.y);
        const double dzm_vz = diff4<M>(c.k, vz_0.y, vz_0.x, vzL.y, vzR.x);
        const double dzp_vx = diff4<M>(c.k, vxR.x, vx_0.y, vx_0.x, vxR.y);
        const double dxm_vz = diff4<M>(c.k, vz_0.y, vz_m1.y, vz_m2.y, vz_p1.y);
        stress_update<M>(c.k, dxp_vx, dzm_vz, dzp_vx, dxm_vz, lm.y, mu.y,
                         dmin2(tr, taper_axis(c.taper_table, c.g.np, z0 + 1, nz)), sxx.y, szz.y, sxz.y);
      }
      txx[i] = sxx;
      tzz[i] = szz;
      txz[i] = sxz;
    }
    __syncthreads();
    // own rows of the new stresses -> global (also the rows the loop above skipped: they carry the old values)
    for (int i = tid; i < (xe - xs) * P2; i += nthreads) {
      const size_t g = (size_t)xs * P2 + i;
      nf2[g] = txx[2 * P2 + i];
      nf3[g] = tzz[2 * P2 + i];
      nf4[g] = txz[2 * P2 + i];
    }
    // 4b. velocities of the own rows, in place
    for (int i = tid; i < (xe - xs) * P2; i += nthreads) {
      const int jr = i / P2, kp = i - jr * P2, x = xs + jr, z0 = 2 * kp;
      const bool rowok = gx0 + x >= 2 && gx0 + x < gnx - 2;
      const bool za = z0 >= 2 && z0 < nz - 2, zb = z0 + 1 >= 2 && z0 + 1 < nz - 2;
      if (!rowok || !(za || zb)) continue;
      const int s = (jr + 2) * P2 + kp;  // stress-tile index of (x, kp)
      const double2 xx_m2 = txx[s - 2 * P2], xx_m1 = txx[s - P2], xx_0 = txx[s], xx_p1 = txx[s + P2];
      const double2 xz_m1 = txz[s - P2], xz_0 = txz[s], xz_p1 = txz[s + P2], xz_p2 = txz[s + 2 * P2];
      const double2 xzL = txz[s - 1], xzR = txz[s + 1];
      const double2 zz_0 = tzz[s], zzL = tzz[s - 1], zzR = tzz[s + 1];
      const double2 b = tb[i];
      const int v = (jr + 4) * P2 + kp;
      double2 vx = tvx[v], vz = tvz[v];
      const double tr = taper_axis(c.taper_table, c.g.np, gx0 + x, gnx);
      if (za) {
        const double dxm_txx = diff4<M>(c.k, xx_0.x, xx_m1.x, xx_m2.x, xx_p1.x);
        const double dzm_txz = diff4<M>(c.k, xz_0.x, xzL.y, xzL.x, xz_0.y);
        const double dxp_txz = diff4<M>(c.k, xz_p1.x, xz_0.x, xz_m1.x, xz_p2.x);
        const double dzp_tzz = diff4<M>(c.k, zz_0.y, zz_0.x, zzL.y, zzR.x);
        velocity_update<M>(c.k, dxm_txx, dzm_txz, dxp_txz, dzp_tzz, b.x,
                           dmin2(tr, taper_axis(c.taper_table, c.g.np, z0, nz)), vx.x, vz.x);
      }
      if (zb) {
        const double dxm_txx = diff4<M>(c.k, xx_0.y, xx_m1.y, xx_m2.y, xx_p1.y);
        const double dzm_txz = diff4<M>(c.k, xz_0.y, xz_0.x, xzL.y, xzR.x);
        const double dxp_txz = diff4<M>(c.k, xz_p1.y, xz_0.y, xz_m1.y, xz_p2.y);
        const double dzp_tzz = diff4<M>(c.k, zzR.x, zz_0.y, zz_0.x, zzR.y);
        velocity_update<M>(c.k, dxm_txx, dzm_txz, dxp_txz, dzp_tzz, b.y,
                           dmin2(tr, taper_axis(c.taper_table, c.g.np, z0 + 1, nz)), vx.y, vz.y);
      }
      tvx[v] = vx;
      tvz[v] = vz;
    }
    // 4c. injections on the new velocities, in list order per target point (rank = position in the point's chain)
    if (inject) {
      __syncthreads();
      double *svx = reinterpret_cast<double *>(tvx), *svz = reinterpret_cast<double *>(tvz);
      const double *sb = reinterpret_cast<const double *>(tb);
      if (fwd) {
        for (int round = 0; round < a.op_rounds; ++round) {
          if (my_rank == round) {
            const double base = __dmul_rn(__dmul_rn(op.weight, inj_a), c.dt);
            const double bb = sb[(op.x - xs) * ld + op.z];
            const double amp = M::kExact ? __ddiv_rn(__dmul_rn(base, bb), op.d4) : (base * bb) * op.rd4;
            double *t = (op.field == 0 ? svx : svz) + (size_t)VROW(op.x) * ld + op.z;
            *t = op.sign > 0 ? __dadd_rn(*t, amp) : __dsub_rn(*t, amp);
          }
          if (round + 1 < a.op_rounds) __syncthreads();
        }
      } else {
        const int x = my_x, z = my_z;
        for (int round = 0; round < a.rec_rounds; ++round) {
          if (my_rank == round) {
            const double bb = sb[(x - xs) * ld + z];
            const double ax = M::kExact ? __ddiv_rn(__dmul_rn(__dmul_rn(c.dt, bb), inj_a), c.dxdz)
                                        : ((c.dt * bb) * inj_a) * c.rdxdz;
            const double az = M::kExact ? __ddiv_rn(__dmul_rn(__dmul_rn(c.dt, bb), inj_b), c.dxdz)
                                        : ((c.dt * bb) * inj_b) * c.rdxdz;
            double *tx = svx + (size_t)VROW(x) * ld + z, *tz = svz + (size_t)VROW(x) * ld + z;
            *tx = __dadd_rn(*tx, ax);
            *tz = __dadd_rn(*tz, az);
          }
          if (round + 1 < a.rec_rounds) __syncthreads();
        }
      }
    }
    __syncthreads();
    // 5. own rows of the new velocities -> global, then publish the step
    {
      double2 *nf0 = reinterpret_cast<double2 *>(m.f[p ^ 1][0]), *nf1 = reinterpret_cast<double2 *>(m.f[p ^ 1][1]);
      for (int i = tid; i < (xe - xs) * P2; i += nthreads) {
        const size_t g = (size_t)xs * P2 + i;
        nf0[g] = tvx[4 * P2 + i];
        nf1[g] = tvz[4 * P2 + i];
      }
    }
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      st_release_u32(a.flags + 32 * (size_t)(blockIdx.y * gridDim.x + blockIdx.x), (unsigned)(step + 1));
    }
  }
#undef VROW
#undef SROW
}

}  // namespace psv
