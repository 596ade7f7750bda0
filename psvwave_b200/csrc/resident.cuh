// Resident whole-loop kernel: ONE launch per forward_simulate / adjoint_simulate call, for grids small enough that a
// time step is shorter than the gap between two kernel launches (the reference's default 250 x 150 configuration:
// loops src/fdModel.cpp:552-751 and :814-958).
//
// The grid is cut into strips of `rows` consecutive rows, one CTA per strip, all CTAs co-resident (cooperative
// launch).  A CTA runs the WHOLE time loop for its strip out of shared memory:
//   * resident for the whole loop: vx vz of rows [xs-3, xe+3), the stress components the own velocities read --
//     txx of rows [xs-2, xe], txz of rows [xs-1, xe+1], tzz of the own rows -- lm mu of those rows, 1/rho of the own
//     rows, the Cerjan taper by row and by column;
//   * per time step those stress rows are updated in place (the halo rows are recomputed redundantly, like the
//     strips of fused.cuh: same explicitly rounded arithmetic on the same inputs, so they equal the owner's values
//     bit for bit and never travel), then -- after a block barrier -- the velocities of the own rows; the thread that
//     owns a point also applies the injections that hit it (src/fdModel.cpp:681-751, :951-958; per-point chains in
//     list order) and sends the new value on its way at once;
//   * the ONLY data exchanged between strips are the new velocities of rows within 3 of a strip edge: every double
//     is written as one 16-byte packet {low word, flag, high word, flag} with flag = number of finished steps (the
//     LL protocol of NCCL), and the neighbours that need the row poll the packets themselves.  No separate
//     counters, no memory fences, no grid-wide barrier: a step costs one L2 round trip of latency.  Packets are
//     double-buffered by step parity; a strip overwrites buffer p only after it has consumed its neighbours' step
//     that followed their last read of p;
//   * snapshot store (:555-572), correlation with the stored forward field (:817-861) and receiver sampling
//     (:576-596) of the NEXT step act on the own rows right after they are published, i.e. while the neighbours'
//     packets are in flight; the final state is written to the ordinary state arrays when the loop ends.
// The arithmetic is the shared point-wise code of stencil_math.cuh in the same order as every other variant, so
// "exact" results are bit-identical to the reference.
#pragma once

#include "fused.cuh"

namespace psv {

constexpr unsigned kResidentSpinLimit = 1u << 22;

__device__ __forceinline__ uint4 ld_packet(const uint4 *p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void st_packet(uint4 *p, double d, unsigned flag) {
  const unsigned lo = (unsigned)__double2loint(d), hi = (unsigned)__double2hiint(d);
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(lo), "r"(flag), "r"(hi), "r"(flag)
               : "memory");
}

template <class M>
__global__ void __launch_bounds__(kResidentThreads, 1) resident_loop_kernel(const ResidentArgs a) {
  extern __shared__ __align__(128) unsigned char psv_smem[];
  double *sm = reinterpret_cast<double *>(psv_smem);
  __shared__ int s_abort;
  // injection operations of the strip: per-point chains (list order), static data, this step's amplitudes
  __shared__ int s_next[kResidentMaxOps], s_meta[kResidentMaxOps];  // meta: bit 0 = z parity, 1 = vz, 2 = subtract
  __shared__ double s_amp[kResidentMaxOps], s_amp2[kResidentMaxOps], s_d4[kResidentMaxOps], s_rd4[kResidentMaxOps];
  const FusedCommon &c = a.c;
  const FusedModel &m = a.models[blockIdx.y];
  const int tid = threadIdx.x, nthreads = blockDim.x;
  const int ld = c.g.ld, P2 = ld >> 1, nz = c.g.nz, nx = c.g.nx;
  const int gx0 = c.g.x_off, gnx = c.g.nx_global;
  const int R = a.rows;
  const int xs = blockIdx.x * R, xe = min(xs + R, nx), own = xe - xs;
  const ResidentLayout L(ld, R);
  double2 *tvx = reinterpret_cast<double2 *>(sm + L.vx), *tvz = reinterpret_cast<double2 *>(sm + L.vz);
  double2 *txx = reinterpret_cast<double2 *>(sm + L.xx), *tzz = reinterpret_cast<double2 *>(sm + L.zz);
  double2 *txz = reinterpret_cast<double2 *>(sm + L.xz);
  double2 *tlm = reinterpret_cast<double2 *>(sm + L.lm), *tmu = reinterpret_cast<double2 *>(sm + L.mu);
  double2 *tb = reinterpret_cast<double2 *>(sm + L.b);
  double *ttz = sm + L.tz, *ttr = sm + L.tr;
  int *head = reinterpret_cast<int *>(sm + L.head);  // first operation of the chain of own pair (jr, kp), or -1
  // tile row of grid row j: velocities (j - xs + 4), stresses / lm / mu (j - xs + 2), b (j - xs)
#define VROW(j) ((j) - xs + 4)
#define SROW(j) ((j) - xs + 2)

  if (tid == 0) s_abort = 0;
  // state S_0 = 0 everywhere; medium and taper of the strip
  for (int i = tid; i < L.lm - L.vx; i += nthreads) sm[L.vx + i] = 0.0;
  {
    const double2 *glm = reinterpret_cast<const double2 *>(m.lm) + (long long)(xs - 2) * P2;
    const double2 *gmu = reinterpret_cast<const double2 *>(m.mu) + (long long)(xs - 2) * P2;
    const double2 *gb = reinterpret_cast<const double2 *>(m.b) + (long long)xs * P2;
    for (int i = tid; i < (own + 4) * P2; i += nthreads) {
      tlm[i] = __ldg(glm + i);
      tmu[i] = __ldg(gmu + i);
    }
    for (int i = tid; i < own * P2; i += nthreads) {
      tb[i] = __ldg(gb + i);
      head[i] = -1;
    }
    for (int i = tid; i < ld; i += nthreads) ttz[i] = taper_axis(c.taper_table, c.g.np, i, nz);
    for (int i = tid; i < own + 4; i += nthreads) ttr[i] = taper_axis(c.taper_table, c.g.np, gx0 + xs - 2 + i, gnx);
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
  // injections: forward = this shot's moment-tensor operations, adjoint = the receivers (adjoint sources)
  const int n_inj = fwd ? oq1 - oq0 : (c.point_mode == POINT_ADJOINT ? rq1 - rq0 : 0);
  double my_weight = 0.0;  // forward: moment entry of operation `tid`
  size_t my_trace = 0;     // row of stf (forward) / of the adjoint sources (adjoint) of injection `tid`
  int my_item = -1;
  if (tid < n_inj) {
    int x, z, meta;
    if (fwd) {
      const RowOp op = c.ops[oq0 + tid];
      x = op.x;
      z = op.z;
      meta = (z & 1) | (op.field ? 2 : 0) | (op.sign > 0 ? 0 : 4);
      my_weight = op.weight;
      my_trace = (size_t)op.source * c.nt;
      s_d4[tid] = op.d4;
      s_rd4[tid] = op.rd4;
    } else {
      x = c.rec_x[rq0 + tid];
      z = c.rec_z[rq0 + tid];
      meta = z & 1;
      my_trace = (size_t)c.rec_id[rq0 + tid] * c.nt;
    }
    s_meta[tid] = meta;
    my_item = (x - xs) * P2 + (z >> 1);
  }
  __syncthreads();
  // chains in list order: walk the list backwards, every operation becomes the head of its point's chain
  for (int q = n_inj - 1; q >= 0; --q) {
    if (tid == q) {
      s_next[q] = head[my_item];
      head[my_item] = q;
    }
    __syncthreads();
  }
  // receiver sampled by this thread (forward): position and trace, constant over the loop
  int smp_v = -1;
  size_t smp_trace = 0;
  if (fwd && c.point_mode == POINT_FORWARD && rq0 + tid < rq1) {
    smp_v = VROW(c.rec_x[rq0 + tid]) * ld + c.rec_z[rq0 + tid];
    smp_trace = (size_t)c.rec_id[rq0 + tid] * c.nt;
  }
  // halo rows of the velocities this strip reads from its neighbours: [xs-3, xs) and [xe, xe+3), inside the grid
  const int lo0 = max(xs - 3, 0), hi1 = min(xe + 3, nx);
  const int n_lo = xs - lo0, n_hi = hi1 - xe;       // rows below / above
  const int halo_packets = 2 * (n_lo + n_hi) * nz;  // vx and vz
  const size_t field_packets = (size_t)nx * ld;     // packets of one field in one parity buffer
  // packet `idx` of the halo: offset inside a parity buffer and the tile element it lands in
  auto locate = [&](int idx, int &src, int &dst) {
    const int rs = idx / nz, col = idx - rs * nz;  // row slot: [vx lo | vx hi | vz lo | vz hi]
    const int per = n_lo + n_hi, f = rs >= per, r = rs - f * per;
    const int j = r < n_lo ? lo0 + r : xe + (r - n_lo);
    src = (int)(f * field_packets) + j * ld + col;
    dst = (f ? L.vz : L.vx) + VROW(j) * ld + col;
  };
  constexpr int U = 4;  // packets a thread keeps in flight; their places are fixed for the whole loop
  int p_src[U], p_dst[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    p_src[u] = p_dst[u] = -1;
    if (tid + u * nthreads < halo_packets) locate(tid + u * nthreads, p_src[u], p_dst[u]);
  }

  // optional phase timing (option "resident_timing"): thread 0 accumulates clock64 deltas per phase
  __shared__ long long t_acc[8];
  if (tid < 8) t_acc[tid] = 0;
  long long t_last = clock64();
#define TICK(k)                      \
  if (a.timing && tid == 0) {        \
    const long long now = clock64(); \
    t_acc[k] += now - t_last;        \
    t_last = now;                    \
  }

  // Side effects of time index `it` on the state BEFORE that step's update, own rows only.
  auto side_effects = [&](int it) {
    if (it % a.si == 0 && (fwd ? a.store != 0 : true)) {
      const size_t snap_off = (size_t)(it / a.si) * a.grid_elems;
      if (fwd) {
        double2 *s = reinterpret_cast<double2 *>(m.snap + snap_off);
        const size_t fs2 = c.snap_field_stride >> 1;
        for (int i = tid; i < own * P2; i += nthreads) {
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
    }
    if (smp_v >= 0) {  // receiver increments from the old velocities (the running sum follows the loop)
      const double pvx = sm[L.vx + smp_v], pvz = sm[L.vz + smp_v];
      m.ux[smp_trace + it] = M::kExact ? __ddiv_rn(__dmul_rn(c.dt, pvx), c.dxdz) : (c.dt * pvx) * c.rdxdz;
      m.uz[smp_trace + it] = M::kExact ? __ddiv_rn(__dmul_rn(c.dt, pvz), c.dxdz) : (c.dt * pvz) * c.rdxdz;
    }
  };

  __syncthreads();
  // What a thread does in the two update phases never changes over the time loop: tile indices, frame masks and taper
  // values of its first item of each phase ride in registers (no integer divisions, table lookups or mask logic per
  // step: three quarters of the instructions of a step were such overhead); further items of wide strips are worked
  // out on the fly.
  struct StressItem {
    int v;           // velocity-tile index of (j, kp)
    unsigned flags;  // 1 update, 2 lower column inside the frame, 4 upper column, 8 txx, 16 tzz, 32 txz
    double tpa, tpb;
  };
  auto stress_item = [&](int i) {
    StressItem it;
    const int jr = i / P2, kp = i - jr * P2, j = xs - 2 + jr, z0 = 2 * kp;
    const bool rowok = gx0 + j >= 2 && gx0 + j < gnx - 2;
    const bool za = z0 >= 2 && z0 < nz - 2, zb = z0 + 1 >= 2 && z0 + 1 < nz - 2;
    it.flags = (rowok && (za || zb) ? 1u : 0u) | (za ? 2u : 0u) | (zb ? 4u : 0u) | (jr <= own + 2 ? 8u : 0u) |
               (jr >= 2 && jr < own + 2 ? 16u : 0u) | (jr >= 1 ? 32u : 0u);
    it.v = (jr + 2) * P2 + kp;
    const double tr = ttr[jr];
    it.tpa = dmin2(tr, ttz[z0]);
    it.tpb = dmin2(tr, ttz[z0 + 1]);
    return it;
  };
  struct VelItem {
    int i, pk;       // own-pair index (velocity tile: i + 4 P2, stress tile: i + 2 P2), packet offset in a parity buffer
    unsigned flags;  // 1 update, 2 lower column, 4 upper column, 8 field is vz, 16 row is read by a neighbour, 32 upper
    int pad;         // column inside the grid
    double tpa, tpb;
  };
  auto vel_item = [&](int w) {
    VelItem it;
    const int pairs = own * P2;
    const int fz = w >= pairs, i = w - fz * pairs;
    const int jr = i / P2, kp = i - jr * P2, x = xs + jr, z0 = 2 * kp;
    const bool rowok = gx0 + x >= 2 && gx0 + x < gnx - 2;
    const bool za = z0 >= 2 && z0 < nz - 2, zb = z0 + 1 >= 2 && z0 + 1 < nz - 2;
    it.i = i;
    it.pad = 0;
    it.pk = (int)(fz * field_packets) + x * ld + z0;
    it.flags = (rowok && (za || zb) ? 1u : 0u) | (za ? 2u : 0u) | (zb ? 4u : 0u) | (fz ? 8u : 0u) |
               ((jr < 3 || jr >= own - 3) && z0 < nz ? 16u : 0u) | (z0 + 1 < nz ? 32u : 0u);
    const double tr = ttr[jr + 2];
    it.tpa = dmin2(tr, ttz[z0]);
    it.tpb = dmin2(tr, ttz[z0 + 1]);
    return it;
  };
  const int n_stress = (own + 4) * P2, n_vel = 2 * own * P2;
  // one record per thread and phase in shared memory when they fit next to the tiles (in registers they spill;
  // same-box A/B: profiles/r2c_resident_kernel.md), else every step works them out again
  const bool cached = L.cached;
  StressItem *s_first = reinterpret_cast<StressItem *>(sm + L.items);
  VelItem *v_firsts = reinterpret_cast<VelItem *>(sm + L.items + 3 * kResidentThreads);
  if (cached) {
    if (tid < n_stress) s_first[tid] = stress_item(tid);
    if (tid < n_vel) v_firsts[tid] = vel_item(tid);
  }
  side_effects(fwd ? 0 : a.nsteps - 1);

  for (int step = 0; step < a.nsteps; ++step) {
    const int it = fwd ? step : a.nsteps - 1 - step;
    // this step's injection amplitudes: independent of the neighbours, in flight while the halo arrives
    if (tid < n_inj) {
      if (fwd) {
        s_amp[tid] = __dmul_rn(__dmul_rn(my_weight, c.stf[my_trace + it]), c.dt);
      } else {
        s_amp[tid] = m.ax[my_trace + it];
        s_amp2[tid] = m.az[my_trace + it];
      }
    }
    // 1. the neighbours' velocities after step-1 (state S_step): poll the packets themselves
    if (step > 0) {
      const uint4 *pk = a.packets + (size_t)(step & 1) * 2 * field_packets;
      const unsigned want = a.flag_base + (unsigned)step;
      auto settle = [&](const uint4 *src, uint4 v, double *dst) {
        unsigned spins = 0;
        while (v.y != want || v.w != want) {
          if (++spins > kResidentSpinLimit) {
            *reinterpret_cast<volatile unsigned *>(a.abort) = 1u;
            s_abort = 1;
            break;
          }
          v = ld_packet(src);
        }
        *dst = __hiloint2double((int)v.z, (int)v.x);
      };
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (p_src[u] >= 0) v[u] = ld_packet(pk + p_src[u]);
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (p_src[u] >= 0) settle(pk + p_src[u], v[u], sm + p_dst[u]);
      for (int idx = tid + U * nthreads; idx < halo_packets; idx += nthreads) {  // wide grids: the rest, one by one
        int src, dst;
        locate(idx, src, dst);
        settle(pk + src, ld_packet(pk + src), sm + dst);
      }
    }
    __syncthreads();
    if (s_abort) return;
    TICK(0);
    // 2. the stress components the own velocities read, in place: txx rows [xs-2, xe], tzz own rows, txz rows
    //    [xs-1, xe+1]
    for (int i = tid; i < n_stress; i += nthreads) {
      const StressItem it = cached && i == tid ? s_first[tid] : stress_item(i);
      if (!(it.flags & 1u)) continue;  // frame rows / pad pairs keep their values
      const bool za = it.flags & 2u, zb = it.flags & 4u;
      const bool do_xx = it.flags & 8u, do_zz = it.flags & 16u, do_xz = it.flags & 32u;
      const int v = it.v;
      const double tpa = it.tpa, tpb = it.tpb;
      const double2 lm = tlm[i], mu = tmu[i];
      if (do_xx) {  // D+x vx and D-z vz feed txx and tzz
        const double2 vx_m1 = tvx[v - P2], vx_0 = tvx[v], vx_p1 = tvx[v + P2], vx_p2 = tvx[v + 2 * P2];
        const double2 vz_0 = tvz[v], vzL = tvz[v - 1], vzR = tvz[v + 1];
        double2 sxx = txx[i];
        double2 szz = do_zz ? tzz[i] : make_double2(0.0, 0.0);
        if (za) {
          const double dxp_vx = diff4<M>(c.k, vx_p1.x, vx_0.x, vx_m1.x, vx_p2.x);
          const double dzm_vz = diff4<M>(c.k, vz_0.x, vzL.y, vzL.x, vz_0.y);
          sxx.x = stress_txx<M>(c.k, dxp_vx, dzm_vz, lm.x, mu.x, tpa, sxx.x);
          if (do_zz) szz.x = stress_tzz<M>(c.k, dxp_vx, dzm_vz, lm.x, mu.x, tpa, szz.x);
        }
        if (zb) {
          const double dxp_vx = diff4<M>(c.k, vx_p1.y, vx_0.y, vx_m1.y, vx_p2.y);
          const double dzm_vz = diff4<M>(c.k, vz_0.y, vz_0.x, vzL.y, vzR.x);
          sxx.y = stress_txx<M>(c.k, dxp_vx, dzm_vz, lm.y, mu.y, tpb, sxx.y);
          if (do_zz) szz.y = stress_tzz<M>(c.k, dxp_vx, dzm_vz, lm.y, mu.y, tpb, szz.y);
        }
        txx[i] = sxx;
        if (do_zz) tzz[i] = szz;
      }
      if (do_xz) {  // D+z vx and D-x vz feed txz
        const double2 vx_0 = tvx[v], vxL = tvx[v - 1], vxR = tvx[v + 1];
        const double2 vz_m2 = tvz[v - 2 * P2], vz_m1 = tvz[v - P2], vz_0 = tvz[v], vz_p1 = tvz[v + P2];
        double2 sxz = txz[i];
        if (za) {
          const double dzp_vx = diff4<M>(c.k, vx_0.y, vx_0.x, vxL.y, vxR.x);
          const double dxm_vz = diff4<M>(c.k, vz_0.x, vz_m1.x, vz_m2.x, vz_p1.x);
          sxz.x = stress_txz<M>(c.k, dzp_vx, dxm_vz, mu.x, tpa, sxz.x);
        }
        if (zb) {
          const double dzp_vx = diff4<M>(c.k, vxR.x, vx_0.y, vx_0.x, vxR.y);
          const double dxm_vz = diff4<M>(c.k, vz_0.y, vz_m1.y, vz_m2.y, vz_p1.y);
          sxz.y = stress_txz<M>(c.k, dzp_vx, dxm_vz, mu.y, tpb, sxz.y);
        }
        txz[i] = sxz;
      }
    }
    __syncthreads();
    TICK(1);
    // 3. velocities of the own rows, in place; injections by the owner of the point; new edge rows sent at once
    {
      const bool publish = step + 1 < a.nsteps;
      uint4 *pk = a.packets + (size_t)((step + 1) & 1) * 2 * field_packets;
      const unsigned flag = a.flag_base + (unsigned)(step + 1);
      // one (z-pair, field) per thread -- all vx pairs first, then all vz pairs, so a warp never holds both: the two
      // fields read disjoint stencil inputs, and twice as many warps share this latency-bound phase
      for (int w = tid; w < n_vel; w += nthreads) {
        const VelItem it = cached && w == tid ? v_firsts[tid] : vel_item(w);
        const int i = it.i, v = i + 4 * P2;
        const bool fz = it.flags & 8u, za = it.flags & 2u, zb = it.flags & 4u;
        double2 *tile = fz ? tvz : tvx;
        double2 val = tile[v];
        const double2 b = tb[i];
        if (it.flags & 1u) {
          const int s = i + 2 * P2;  // stress-tile index of (x, kp)
          const double tpa = it.tpa, tpb = it.tpb;
          if (!fz) {  // vx: D-x txx + D-z txz
            const double2 xx_m2 = txx[s - 2 * P2], xx_m1 = txx[s - P2], xx_0 = txx[s], xx_p1 = txx[s + P2];
            const double2 xz_0 = txz[s], xzL = txz[s - 1], xzR = txz[s + 1];
            if (za)
              val.x = velocity_component<M>(c.k, diff4<M>(c.k, xx_0.x, xx_m1.x, xx_m2.x, xx_p1.x),
                                            diff4<M>(c.k, xz_0.x, xzL.y, xzL.x, xz_0.y), b.x, tpa, val.x);
            if (zb)
              val.y = velocity_component<M>(c.k, diff4<M>(c.k, xx_0.y, xx_m1.y, xx_m2.y, xx_p1.y),
                                            diff4<M>(c.k, xz_0.y, xz_0.x, xzL.y, xzR.x), b.y, tpb, val.y);
          } else {  // vz: D+x txz + D+z tzz
            const double2 xz_m1 = txz[s - P2], xz_0 = txz[s], xz_p1 = txz[s + P2], xz_p2 = txz[s + 2 * P2];
            const double2 zz_0 = tzz[s], zzL = tzz[s - 1], zzR = tzz[s + 1];
            if (za)
              val.x = velocity_component<M>(c.k, diff4<M>(c.k, xz_p1.x, xz_0.x, xz_m1.x, xz_p2.x),
                                            diff4<M>(c.k, zz_0.y, zz_0.x, zzL.y, zzR.x), b.x, tpa, val.x);
            if (zb)
              val.y = velocity_component<M>(c.k, diff4<M>(c.k, xz_p1.y, xz_0.y, xz_m1.y, xz_p2.y),
                                            diff4<M>(c.k, zzR.x, zz_0.y, zz_0.x, zzR.y), b.y, tpb, val.y);
          }
        }
        for (int q = head[i]; q >= 0; q = s_next[q]) {  // this pair's injections, in list order; my field's among them
          const int meta = s_meta[q];
          const double bb = (meta & 1) ? b.y : b.x;
          double &t = (meta & 1) ? val.y : val.x;
          if (fwd) {
            if ((((meta >> 1) & 1) != 0) != fz) continue;
            const double amp = M::kExact ? __ddiv_rn(__dmul_rn(s_amp[q], bb), s_d4[q])
                                         : __dmul_rn(__dmul_rn(s_amp[q], bb), s_rd4[q]);
            t = (meta & 4) ? __dsub_rn(t, amp) : __dadd_rn(t, amp);
          } else {
            const double dtb = __dmul_rn(c.dt, bb), src = fz ? s_amp2[q] : s_amp[q];
            t = __dadd_rn(t, M::kExact ? __ddiv_rn(__dmul_rn(dtb, src), c.dxdz) : __dmul_rn(__dmul_rn(dtb, src), c.rdxdz));
          }
        }
        tile[v] = val;
        if (publish && (it.flags & 16u)) {  // rows a neighbour reads (state S_{step+1})
          uint4 *p0 = pk + it.pk;
          st_packet(p0, val.x, flag);
          if (it.flags & 32u) st_packet(p0 + 1, val.y, flag);
        }
      }
    }
    __syncthreads();
    TICK(2);
    // 4. what the next step does to its starting state (snapshot, correlation, receiver samples): own rows, while
    //    the packets travel
    if (step + 1 < a.nsteps) side_effects(fwd ? step + 1 : a.nsteps - 2 - step);
    TICK(3);
  }
  // the final state of the own rows -> the ordinary state arrays (buffer nt & 1, where the per-step kernels leave it)
  {
    const int p = a.nsteps & 1;
    double2 *f0 = reinterpret_cast<double2 *>(m.f[p][0]), *f1 = reinterpret_cast<double2 *>(m.f[p][1]);
    double2 *f2 = reinterpret_cast<double2 *>(m.f[p][2]), *f3 = reinterpret_cast<double2 *>(m.f[p][3]);
    double2 *f4 = reinterpret_cast<double2 *>(m.f[p][4]);
    for (int i = tid; i < own * P2; i += nthreads) {
      const size_t g = (size_t)xs * P2 + i;
      f0[g] = tvx[4 * P2 + i];
      f1[g] = tvz[4 * P2 + i];
      f2[g] = txx[2 * P2 + i];
      f3[g] = tzz[2 * P2 + i];
      f4[g] = txz[2 * P2 + i];
    }
  }
  if (a.timing && tid == 0)
    for (int k = 0; k < 8; ++k) a.timing[8 * (size_t)(blockIdx.y * gridDim.x + blockIdx.x) + k] = t_acc[k];
#undef TICK
#undef VROW
#undef SROW
}

}  // namespace psv
