// Variant 4: TWO time steps per launch (temporal blocking through shared memory).
//
// The state of a P-SV step is 5 words per point and a step is HBM-bound, so the cheapest byte
// is the one never written: a CTA runs step t -> t+1 and step t+1 -> t+2 back to back on the
// same rows and the intermediate state t+1 never leaves the SM.
//     reads  vx vz txx tzz txz (state t) + lm mu b      8 words / point
//     writes vx vz txx tzz txz (state t+2)              5 words / point
// = 104 B per TWO grid-point updates = 52 B per update (variant 3: 104 B, two-sweep: 176 B).
//
// A CTA is a three-stage warp-specialised pipeline along x:
//   producer        TMA bulk copies of the rows of state t into ring T (as variant 3), issued by
//                     lane 0 of stage-A warp 0 between its own rows
//   NW stage-A warps  the march of fused.cuh on ring T; their results -- new stresses of row i,
//                     new velocities of row i-2, and the medium rows stage B will want -- go
//                     into slot i of a second shared-memory ring X instead of global memory
//   NW stage-B warps  the same march, fed from ring X (the rows of state t+1), results to
//                     global memory (state t+2)
// Slots of ring X are guarded by full/empty mbarriers (NW arrivals each way).  Stage B's
// iteration j needs the stresses of slot j, vz of slot j+3 and vx of slot j+4, so a slot lives
// for 5 iterations and the ring holds kPairSlots >= 6.
//
// Geometry.  Stage B owns the CTA's tile [tb0, tb0 + tile_pairs) x [xs, xe).  Stage A must
// deliver state t+1 on B's whole read footprint: its warps own (= compute validly) the
// NW*28 pairs starting at tb0 - 2 and march the rows [xs-4, xe+4) (clamped to the grid), and
// ring T starts another 2 pairs to the left.  Redundant work per CTA: 4 pairs and 8 rows.
//
// Global side effects keep the reference's order: receiver samples u[it] (stage A, from state
// t) and u[it+1] (stage B, from state t+1, after A's release of the slot that carries the
// velocities -- A's global store of u[it] happens-before through the slot's mbarrier),
// snapshot store / kernel correlation of state t (stage A; the host never pairs a step whose
// SECOND half would need one), injections in both stages on every lane that computes the
// target point (stage A's halo copies included, they feed stage B).
#pragma once

#include "fused.cuh"

namespace psv {

constexpr int kPairSlots = 6;

template <int NW>
struct PairGeom {
  static constexpr int kOwnedA = NW * kFusedOwnedPairs;  // pairs stage A computes validly
  static constexpr int kMaxTile = kOwnedA - 4;           // widest tile stage B can own
  static constexpr int kPairsT = kOwnedA + 4;            // ring T row segment: + 2 halo pairs each side
  static constexpr int kSegT = kPairsT * 16;
  static constexpr int kStageT = kSegT * kPipeArrays;
  static constexpr int kSegX = kOwnedA * 16;
  static constexpr int kSlotX = kSegX * kPipeArrays;
  // ring X array order inside a slot: the same as ring T (vx vz txx tzz txz lm mu b)
};

template <int NW>
constexpr size_t pair_smem_bytes(int stages_t) {
  return 128 * ((2 * (size_t)(stages_t + kPairSlots) * 8 + 127) / 128) + (size_t)stages_t * PairGeom<NW>::kStageT +
         (size_t)kPairSlots * PairGeom<NW>::kSlotX;
}

// Sink of stage A: slot (i - jA0) mod kPairSlots of ring X.
struct SlotSink {
  unsigned char *ring;
  unsigned full0, empty0;
  int seg, slot_bytes;
  int my_off;  // byte offset of my pair inside a segment
  int slot;
  unsigned phase;
  int lane;
  bool own;
  bool acquired;
  bool wrapped;  // false during the first pass over the ring: nothing to wait for.  (The usual
                 // "wait for parity 1 on a fresh barrier" shortcut is wrong here: stage B hands
                 // back the slots it skips before stage A ever touches them, which completes
                 // phase 0 early and would leave a parity-1 wait hanging.)
  __device__ __forceinline__ void acquire() {
    if (!acquired) {
      if (wrapped) mbar_wait(empty0 + 8 * slot, phase ^ 1);
      acquired = true;
    }
  }
  __device__ __forceinline__ void stress(int, long long, double2 txx, double2 tzz, double2 txz, double2 lm,
                                         double2 mu) {
    acquire();
    if (own) {
      unsigned char *s = ring + (size_t)slot * slot_bytes + my_off;
      *reinterpret_cast<double2 *>(s + 2 * seg) = txx;
      *reinterpret_cast<double2 *>(s + 3 * seg) = tzz;
      *reinterpret_cast<double2 *>(s + 4 * seg) = txz;
      *reinterpret_cast<double2 *>(s + 5 * seg) = lm;
      *reinterpret_cast<double2 *>(s + 6 * seg) = mu;
    }
  }
  __device__ __forceinline__ void velocity(int, long long, double2 vx, double2 vz, double2 b) {
    if (own) {
      unsigned char *s = ring + (size_t)slot * slot_bytes + my_off;
      *reinterpret_cast<double2 *>(s) = vx;
      *reinterpret_cast<double2 *>(s + seg) = vz;
      *reinterpret_cast<double2 *>(s + 7 * seg) = b;
    }
  }
  __device__ __forceinline__ void commit() {
    __syncwarp();
    if (lane == 0) mbar_arrive(full0 + 8 * slot);
    acquired = false;
    if (++slot == kPairSlots) {
      slot = 0;
      phase ^= 1;
      wrapped = true;
    }
  }
};

// Row source of stage B: ring X.  Stage-A iteration i lives in slot (i - jA0) mod kPairSlots.
struct PairRows {
  const unsigned char *ring;
  unsigned full0, empty0;
  int seg, slot_bytes;
  int lane_off;      // byte offset of my pair inside a segment
  int jA0, jA_last;  // first / last stage-A iteration
  int xsA, xeA;      // rows whose velocities stage A delivers
  int lane;
  __device__ __forceinline__ int slot_of(int i) const { return (i - jA0) % kPairSlots; }
  __device__ __forceinline__ double2 lift(int i, int array) const {
    return *reinterpret_cast<const double2 *>(ring + (size_t)slot_of(i) * slot_bytes + array * seg + lane_off);
  }
  // velocities of row `row` (array 0 = vx, 1 = vz); rows stage A does not deliver lie outside the
  // grid and read as the guard band's zeros
  __device__ __forceinline__ double2 vel(int row, int array) const {
    if (row < xsA || row >= xeA) return make_double2(0.0, 0.0);
    return lift(row + 2, array);
  }
  // Parity waits can only tell "phase n" from "phase n+1", so the slots are waited for strictly in
  // stage-A order, one phase after the other (a wait on a slot that is already full costs one
  // try_wait); `waited` is the next stage-A iteration not yet known to be complete.
  int waited, w_slot;
  unsigned w_phase;
  __device__ __forceinline__ void wait_upto(int i) {
    i = min(i, jA_last);
    while (waited <= i) {
      mbar_wait(full0 + 8 * w_slot, w_phase);
      ++waited;
      if (++w_slot == kPairSlots) {
        w_slot = 0;
        w_phase ^= 1;
      }
    }
  }
  __device__ __forceinline__ void prologue(int j, long long, MarchWindow &w) {
    // slots of stage-A iterations before j carry nothing stage B reads: hand them back at once
    if (lane == 0)
      for (int i = jA0; i < j; ++i) mbar_arrive(empty0 + 8 * slot_of(i));
    wait_upto(j + 3);  // vx row j+1
    w.vx_m2 = vel(j - 2, 0);
    w.vx_m1 = vel(j - 1, 0);
    w.vx_0 = vel(j, 0);
    w.vx_p1 = vel(j + 1, 0);
    w.vz_m2 = vel(j - 2, 1);
    w.vz_m1 = vel(j - 1, 1);
    w.vz_0 = vel(j, 1);
  }
  __device__ __forceinline__ void fetch(int j, long long, RowData &d) {
    wait_upto(j + 4);
    d.vx_p2 = vel(j + 2, 0);
    d.vz_p1 = vel(j + 1, 1);
    d.txx = lift(j, 2);
    d.tzz = lift(j, 3);
    d.txz = lift(j, 4);
    d.lm = lift(j, 5);
    d.mu = lift(j, 6);
    d.b = lift(j, 7);
    cur = j;
  }
  // stage B runs plain steps only (the host never pairs a step whose second half correlates)
  __device__ __forceinline__ void fetch_corr(const FusedCommon &, const FusedModel &, long long, bool, CorrData &) const {}
  __device__ __forceinline__ void release() {
    __syncwarp();
    if (lane == 0) mbar_arrive(empty0 + 8 * slot_of(cur));
  }
  int cur;
};

// Row source of stage A: ring T as in variant 3, but without a producer warp -- 2*NW warps fill
// the four SM sub-partitions evenly (3 each for NW = 6), which is what leaves 168 registers per
// thread; a 13th warp would cap everyone at 128.  Lane 0 of stage-A warp 0 issues the bulk
// copies: before it consumes row j it makes sure the rows up to j + stages - 1 are in flight.
template <int SEG>
struct SelfFedRows {
  PipeRows ring;
  const FusedModel *m;
  const double *f0;      // old vx; old field k = f0 + k * field_stride
  size_t field_stride;
  long long ld2n;
  int kp_t;              // first pair of a row segment
  int next, last;        // next row to issue, last row of the march
  bool producer;
  int pslot;
  unsigned pphase;
  __device__ __forceinline__ void issue(int j) {
    const unsigned bar = ring.full0 + 8 * pslot;
    mbar_wait(ring.empty0 + 8 * pslot, pphase ^ 1);
    mbar_expect_tx(bar, SEG * kPipeArrays);
    const unsigned dst = smem_u32(ring.ring) + pslot * (SEG * kPipeArrays);
    const double2 *base = reinterpret_cast<const double2 *>(f0) + (long long)j * ld2n + kp_t;
    const long long fs2 = (long long)(field_stride >> 1);
    bulk_g2s(dst, base + 2 * ld2n, SEG, bar);
    bulk_g2s(dst + SEG, base + fs2 + ld2n, SEG, bar);
    bulk_g2s(dst + 2 * SEG, base + 2 * fs2, SEG, bar);
    bulk_g2s(dst + 3 * SEG, base + 3 * fs2, SEG, bar);
    bulk_g2s(dst + 4 * SEG, base + 4 * fs2, SEG, bar);
    const long long off = (long long)j * ld2n + kp_t;
    bulk_g2s(dst + 5 * SEG, reinterpret_cast<const double2 *>(m->lm) + off, SEG, bar);
    bulk_g2s(dst + 6 * SEG, reinterpret_cast<const double2 *>(m->mu) + off, SEG, bar);
    bulk_g2s(dst + 7 * SEG, reinterpret_cast<const double2 *>(m->b) + off - 2 * ld2n, SEG, bar);
    if (++pslot == ring.stages) {
      pslot = 0;
      pphase ^= 1;
    }
  }
  __device__ __forceinline__ void feed(int upto) {
    if (producer) {
      upto = min(upto, last);
      while (next <= upto) issue(next++);
    }
    __syncwarp();
  }
  __device__ __forceinline__ void prologue(int j, long long r, MarchWindow &w) {
    next = j;
    feed(j + ring.stages - 2);  // keep one slot back: the consumers have released nothing yet
    ring.prologue(j, r, w);
  }
  __device__ __forceinline__ void fetch(int j, long long r, RowData &d) {
    ring.fetch(j, r, d);  // consumes row j
    cur = j;
  }
  __device__ __forceinline__ void fetch_corr(const FusedCommon &c, const FusedModel &fm, long long r, bool lane_ok,
                                             CorrData &q) const {
    ring.fetch_corr(c, fm, r, lane_ok, q);  // ring T carries no correlation segments: straight from global memory
  }
  __device__ __forceinline__ void release() {
    ring.release();               // this warp's share of the slot
    feed(cur + ring.stages - 1);  // refill; blocks until the slowest stage-A warp has left row cur - 1
  }
  int cur;
};

template <class M, int MODE, int NW>
__global__ void __launch_bounds__(2 * NW * 32) fused_pair_kernel(const FusedCommon c,
                                                                       const FusedModel *__restrict__ models,
                                                                       const int stages) {
  using G = PairGeom<NW>;
  extern __shared__ __align__(128) unsigned char psv_smem[];
  unsigned long long *bars = reinterpret_cast<unsigned long long *>(psv_smem);
  unsigned char *ring_t = psv_smem + 128 * ((2 * (stages + kPairSlots) * 8 + 127) / 128);
  unsigned char *ring_x = ring_t + (size_t)stages * G::kStageT;
  const unsigned full_t = smem_u32(bars), empty_t = full_t + 8 * stages;
  const unsigned full_x = empty_t + 8 * stages, empty_x = full_x + 8 * kPairSlots;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(full_t + 8 * s, 1);
      mbar_init(empty_t + 8 * s, NW);
    }
    for (int s = 0; s < kPairSlots; ++s) {
      mbar_init(full_x + 8 * s, NW);
      mbar_init(empty_x + 8 * s, NW);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  const FusedModel &m = models[blockIdx.z];
  const int nx = c.g.nx;
  const long long ld2n = c.g.ld >> 1;
  const int xs = blockIdx.y * c.strip;
  const int xe = min(xs + c.strip, nx);
  const int xsA = max(xs - 4, 0), xeA = min(xe + 4, nx);
  const int tb0 = blockIdx.x * c.tile_pairs;  // first pair of stage B's tile
  const int ta0 = tb0 - 2;                    // first pair stage A owns
  const int kp_t = ta0 - 2;                   // first pair of ring T's row segments

  if (warp < NW) {
    // ---- stage A: state t -> t+1, ring T -> ring X ----
    PipeRows ring{ring_t, full_t, empty_t, stages, G::kSegT, G::kStageT, (warp * kFusedOwnedPairs + lane) * 16,
                  0,      0u,     lane,    m.f[c.parity][0], m.f[c.parity][1], ld2n, false};
    SelfFedRows<G::kSegT> rows{ring, &m, m.f[c.parity][0], c.field_stride, ld2n, kp_t, xsA - 2, xeA + 1,
                               warp == 0 && lane == 0, 0, 0u, 0};
    const int kp = kp_t + warp * kFusedOwnedPairs + lane;
    const bool own = lane >= 2 && lane < 30;
    const bool mine = own && kp >= tb0 && kp < tb0 + c.tile_pairs && kp < (int)ld2n;
    SlotSink sink{ring_x, full_x, empty_x, G::kSegX, G::kSlotX, (warp * kFusedOwnedPairs + lane - 2) * 16,
                  0,      0u,     lane,    own,      false, false};
    fused_march<M, MODE>(c, m, rows, sink, lane, kp, MarchRange{xsA, xeA, xs, xe, own, mine, c.it});
    return;
  }

  // ---- stage B: state t+1 -> t+2, ring X -> global ----
  {
    const int wb = warp - NW;
    const int idx = wb * kFusedOwnedPairs + lane;  // position inside stage A's owned range
    const bool inside = idx < G::kOwnedA;
    PairRows rows{ring_x, full_x, empty_x, G::kSegX, G::kSlotX, (inside ? idx : 0) * 16,
                  xsA - 2, xeA + 1, xsA,   xeA,      lane,     xsA - 2, 0, 0u, 0};
    const int kp = ta0 + idx;
    const bool own = lane >= 2 && lane < 30 && kp >= tb0 && kp < tb0 + c.tile_pairs && kp < (int)ld2n;
    GlobalSink sink{m.f[c.parity ^ 1][0], m.f[c.parity ^ 1][1], m.f[c.parity ^ 1][2], m.f[c.parity ^ 1][3],
                    m.f[c.parity ^ 1][4], xs, xe, own};
    fused_march<M, MODE_PLAIN>(c, m, rows, sink, lane, kp, MarchRange{xs, xe, xs, xe, own, own, c.it + c.it_step});
  }
}

}  // namespace psv
