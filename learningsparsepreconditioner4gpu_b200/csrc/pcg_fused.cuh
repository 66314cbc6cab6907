// Whole PCG solve of a LARGE system in ONE persistent kernel (driver "mode 4"), single GPU or row-partitioned over the
// GPUs of an NVSwitch box.  It replaces the 5-launch iteration of pcg.cu / the 8-launch iteration of dist.cu:
//
//   P1   t = G^T r [/ d]                                    (matrix sweep, gathers r)
//   --- barrier B1: local grid barrier; row-partitioned: + the neighbours' "boundary rows of t stored" flags (no all-reduce)
//   P2   z = G t + eps r [/ d] ; rho = r.z                  (matrix sweep, gathers t)
//   --- barrier B2 (cross-rank): rho -> beta, and ||r||^2 of the update made at the end of the previous iteration ->
//       iteration count, history, convergence / max_iter test (the P1/P2 sweeps that ran ahead of a converged test only
//       wrote t and z)
//   V2   p = z + beta p                                     (vector phase, owned rows + halo)
//   --- local grid barrier
//   P3   q = A p ; p.q                                      (matrix sweep, gathers p)
//   --- barrier B3 (cross-rank): p.q -> alpha (breakdown if <= 0)
//   V1   x += alpha p ; r -= alpha q ; ||r||^2              (vector phase, owned rows + halo of r)
//   --- local grid barrier
//
// Nothing is launched, captured or polled by the host while the solve runs.  (A first version folded V1 and V2 into
// the sweeps by recomputing r' = r - alpha q and p' = z + beta p inside the gathers: two gathers per matrix entry made
// the sweeps 30-90 % slower than the vector phases they saved -- measured 1.07 vs 0.84 ms per iteration at 216^3 -- so
// the updates are separate phases again; at 8 ranks the vectors of a rank live in L2 and these phases cost ~10 us.)
//
// Each sweep is the TMA-staged streaming SpMV of spmv.cuh (row blocks staged by one producer lane with cp.async.bulk +
// mbarrier, consumer warps row-parallel out of shared memory, lossless fp32-value / 16-bit-column-step streams, row
// sums in scipy's sequential rounded order).  The matrices never change, so the producer simply runs ahead through the
// cyclic schedule G^T, G, A, G^T, ... : the first blocks of the next sweep are already in shared memory while the CTA
// waits at a barrier or runs a vector phase (no pipeline ramp per sweep, no launch tail).
//
// Cross-rank barrier = monotonic arrive counter + release word in global memory (cooperative launch guarantees
// co-residency).  Every CTA deposits its two partial sums before arriving; warp 0 of CTA 0 sums them in CTA order
// (deterministic), and all CTAs read the totals that travel with the release -> identical scalars and identical
// branches everywhere.  Gathered vectors are read with plain ld.global (L1-cached); the acquire that ends every barrier
// invalidates L1, which is what makes re-reading vectors other SMs rewrote legal (ld.global.nc would not be).
// The local barrier is the same counter scheme without values or master.
//
// Row-partitioned (world > 1): every rank runs this kernel on its own rows.  The thread that computes an owned BOUNDARY
// entry of t, z or q also stores it straight into the neighbours' extended vectors (peer stores over NVLink, "dual
// store"), fences at system scope and arrives; CTA 0's warp then all-reduces the two totals through every rank's
// peer-mapped exchange block (value + epoch per (barrier kind, source rank), summed in RANK order -> identical bits on
// all ranks) before it releases the local grid.  The release/acquire chain (peer store -> fence.sys -> arrive ->
// master acquire -> st.release.sys epoch -> peer's acquire -> peer's release of its grid) also orders the halo data,
// so one barrier = halo exchange + all-reduce + grid sync.  Halo entries of r and p are updated locally by the vector
// phases from the halos of q and z: they carry the bits their owner computes and are never sent.
#pragma once
#include "pcg_kernels.cuh"

namespace b200spai {

constexpr int FUSED_MAXR = 16;
enum { FB_INIT = 0, FB_RR = 1, FB_RHO = 2, FB_PQ = 3, FB_READY = 4, FB_HALO = 5 /* +0 t, +1 z, +2 q */, FB_KINDS = 8 };

struct FusedMat {
  const int* rowptr;
  const int* col;
  const double* val;    // canonical arrays (rows longer than a stage)
  const void* sval;     // streamed values: float or double
  const void* scol;     // streamed columns: u16 steps or int
  const int* rowfirst;  // first column of each row (u16 stream only)
  const int4* blk;      // row blocks {r0, rowptr[r0], r1, rowptr[r1]}, any order
  int nblk;
  int nbnd;             // (row-partitioned) the first nbnd blocks hold the rows a neighbour's halo needs
};

struct FusedSync {  // one per solver, zeroed before every launch
  unsigned long long arrive;
  unsigned long long pad0[15];
  // release of the value barrier, LL style: three 16-byte pairs {low half | gen, high half | gen} of total 0, total 1 and
  // the ok flag -- the totals travel WITH the release, a waiting CTA needs one L2 round trip, not two
  unsigned long long rel[6];
  unsigned long long pad1[10];
  unsigned long long light;  // arrive counter of the value-less local barrier
  unsigned long long pad2[15];
  unsigned long long light_rel;  // its release word (generation), written by the last arriver
  unsigned long long pad2b[15];
  unsigned long long bdone[3][16];  // [t, z, q][0]: CTAs that finished (and fenced) the boundary blocks of such a sweep
};

struct alignas(16) FusedXchg {  // one per rank, inside the peer-mapped segment; never reset (epochs are monotonic across solves)
  // [kind][source rank] = two words {low 32 bits of the value << 32 | epoch, high 32 bits of the value | epoch}
  unsigned long long w[FB_KINDS][FUSED_MAXR][2];
  // [t, z, q][source rank] = epoch of the last sweep of that kind whose boundary rows that rank has stored into this rank's halo
  unsigned long long hflag[3][FUSED_MAXR];
};

struct FusedPeer {
  int rank, world, n_push, n_recv;
  int push_peer[FUSED_MAXR], recv_peer[FUSED_MAXR];  // ranks whose halo holds my rows / whose rows my halo holds
  int push_lo[FUSED_MAXR], push_hi[FUSED_MAXR];  // owned local rows [lo, hi) a neighbour's extended range contains
  double* push_r[FUSED_MAXR];                    // address of row push_lo[s] inside the neighbour's extended vector
  double* push_t[FUSED_MAXR];
  double* push_z[FUSED_MAXR];
  double* push_q[FUSED_MAXR];
  FusedXchg* xchg[FUSED_MAXR];  // exchange block of every rank (own included)
  unsigned long long* epoch;    // [FB_KINDS] local monotonic counters (device memory, survive across solves)
};

struct FusedArgs {
  FusedMat A, G, GT;
  int n;                 // owned rows
  int off;               // extended index of owned row 0 (= number of low halo entries)
  int hi_halo;           // extended entries behind the owned slice
  const double* b;       // [n]
  double* x;             // [n]
  double *r, *p, *z, *q, *t;  // EXTENDED vectors (index 0 = first low halo entry)
  const double* d;       // [n] diagonal (scaled SPAI) or nullptr
  double eps;
  int precond;
  PcgScalars* sc;
  double* partials;      // [4 * gridDim.x] 8-byte LL words (two tagged pairs per CTA), zeroed before every launch
  double* hist;
  FusedSync* sync;
  long long spin_limit;  // clock64 ticks before a wait gives up
  // optional [9]: ns in P1+B1, P2+B2, V2+barrier, P3+B3, V1+barrier (CTA 0's view), iterations, then (world > 1) ns CTA 0
  // waited for the last local CTA, ns from there to the last peer's contribution, cross-rank barriers timed
  unsigned long long* prof;
  FusedPeer peer;
};

// ---- memory-model helpers -------------------------------------------------------------------------------------
__device__ __forceinline__ double ld_coh(const double* p) {  // coherent (L1-cached) load; NOT ld.global.nc
  double v;
  asm volatile("ld.global.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_gpu_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void red_release_gpu_add_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void f_st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long f_ld_relaxed_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
template <int NT>
__device__ __forceinline__ void bar_consumers() {
  asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
}

// CTA-wide OR over the consumer threads (named barrier 1), doubling as the barrier itself.
template <int NT>
__device__ __forceinline__ bool bar_or_consumers(bool pred) {
  uint32_t r;
  asm volatile(
      "{\n"
      ".reg .pred p, q;\n"
      "setp.ne.u32 q, %1, 0;\n"
      "bar.red.or.pred p, 1, %2, q;\n"
      "selp.u32 %0, 1, 0, p;\n"
      "}\n"
      : "=r"(r)
      : "r"((uint32_t)pred), "n"(NT)
      : "memory");
  return r != 0;
}
__device__ __forceinline__ void st_relaxed_sys_v2(unsigned long long* p, unsigned long long w0, unsigned long long w1) {
  asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(w0), "l"(w1) : "memory");
}
__device__ __forceinline__ void ld_relaxed_sys_v2(const unsigned long long* p, unsigned long long& w0, unsigned long long& w1) {
  asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(p) : "memory");
}

// Shared scratch of the barrier code: red[0..31] warp partials, [32..35] broadcast, then (CTA 0 only) the epoch counters
// of the cross-rank exchange and the barrier timers, all 8-byte slots.
constexpr int FRED_EPOCH = 36;              // FB_KINDS slots
constexpr int FRED_T_LOCAL = 44;            // ns CTA 0 waited for the last local CTA's partials (after its own arrival)
constexpr int FRED_T_PEER = 45;             // ns from there to the last peer's contribution
constexpr int FRED_N_BAR = 46;              // cross-rank barriers timed
constexpr int FRED_SLOTS = 48;
static_assert(FRED_EPOCH + FB_KINDS <= FRED_T_LOCAL, "epoch slots overlap the timers");

// 64 bits as two tagged words {low half | tag, high half | tag} (NCCL-LL style): 8-byte words are single-copy atomic, so a
// reader that finds `tag` in both words holds the value the writer stored with it -- value and flag in one access.
__device__ __forceinline__ void ll_store_gpu(unsigned long long* slot, unsigned long long bits, unsigned long long tag32) {
  asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(slot), "l"(((bits & 0xffffffffull) << 32) | tag32),
               "l"((bits & 0xffffffff00000000ull) | tag32) : "memory");
}
__device__ __forceinline__ bool ll_load_gpu(const unsigned long long* slot, unsigned long long tag32, unsigned long long& bits) {
  unsigned long long w0, w1;
  asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(slot) : "memory");
  bits = (w1 & 0xffffffff00000000ull) | (w0 >> 32);
  return (w0 & 0xffffffffull) == tag32 && (w1 & 0xffffffffull) == tag32;
}

// (row-partitioned) CTA 0, thread 0: wait until every neighbour whose rows this rank's halo holds has flagged the sweep
// `halo_epoch` of kind `hk` (0: t, 1: z, 2: q), then fence at system scope.  The flag is raised by the neighbour's last
// boundary CTA early in ITS sweep (the boundary blocks go first), so this normally returns at once.
__device__ __forceinline__ bool fused_wait_halo(const FusedArgs& a, int hk, unsigned long long halo_epoch) {
  bool ok = true;
  const long long w0 = clock64();
  for (int k = 0; k < a.peer.n_recv && ok; ++k) {
    const unsigned long long* f = &a.peer.xchg[a.peer.rank]->hflag[hk][a.peer.recv_peer[k]];
    while (f_ld_relaxed_sys(f) < halo_epoch)
      if (clock64() - w0 > a.spin_limit) {
        ok = false;
        break;
      }
  }
  asm volatile("fence.acq_rel.sys;" ::: "memory");
  return ok;
}

// Grid barrier + deterministic reduction of (v0, v1) + (world > 1) all-reduce over the ranks.  Called by every
// CONSUMER thread of every CTA; returns false when a wait timed out (peer or CTA gone): the solve is then abandoned.
//
// Local part: every CTA publishes its two partial sums as LL words tagged with the barrier generation (arrival and value
// in one store, after a gpu-scope fence that orders the CTA's vector writes); the consumer threads of CTA 0 poll one
// slot each (CTAs tid, tid + 256, ...), sum in a fixed order (thread-strided, warp tree, warps in index order); warp 0
// exchanges with the other ranks and publishes the totals as LL words every CTA polls -- two L2 round trips end to end.
//
// Cross-rank part: the rank's total travels as an LL pair into slot [kind][source rank] of every rank's exchange block;
// the receiver polls until both words carry the epoch and sums IN RANK ORDER (identical bits on every rank).  `both`:
// v0 under `kind` AND v1 under FB_RR in the same exchange (lanes 0..15 / 16..31).
//
// Halo data.  halo_kind >= 0 (sweep barriers): boundary rows were flagged by fused_boundary_done when they were stored;
// CTA 0 waits for its neighbours' flags (fused_wait_halo) BEFORE it joins the barrier, so the scalar exchange itself
// needs no system-scope fence on either side.  halo_kind < 0 (init / ready barriers): `pushed` CTAs fence at system
// scope before they arrive and the master fences around the exchange, which then also orders the pushed data.
template <int NCW>
__device__ __forceinline__ bool fused_barrier(const FusedArgs& a, unsigned long long& gen, int kind, double v0, double v1,
                                              bool pushed, double* red, double& t0, double& t1, bool both = false,
                                              int halo_kind = -1, unsigned long long halo_epoch = 0ull) {
  constexpr int NCT = NCW * 32;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned G = gridDim.x;
  const bool master = blockIdx.x == 0;
  const bool fenced = halo_kind < 0;  // the exchange itself orders pushed data
  gen += 1;
  const unsigned long long g32 = gen & 0xffffffffull;
  unsigned long long* ll = reinterpret_cast<unsigned long long*>(a.partials);  // [G][4] LL words
  unsigned long long* red64 = reinterpret_cast<unsigned long long*>(red);
  const bool timed = master && a.prof != nullptr && a.peer.world > 1;
  unsigned long long ns_arrive = 0ull;
  if (master && tid == 0) {
    red[35] = 1.0;
    if (!fenced && a.peer.world > 1 && a.peer.n_recv > 0 && !fused_wait_halo(a, halo_kind, halo_epoch)) red[35] = 0.0;
  }
  v0 = warp_sum(v0);
  v1 = warp_sum(v1);
  if (lane == 0) {
    red[warp] = v0;
    red[16 + warp] = v1;
  }
  const bool cta_pushed = bar_or_consumers<NCT>(pushed);
  if (warp == 0) {
    double c0 = (lane < NCW) ? red[lane] : 0.0, c1 = (lane < NCW) ? red[16 + lane] : 0.0;
    c0 = warp_sum(c0);
    c1 = warp_sum(c1);
    if (lane < 2) {
      if (fenced && cta_pushed) asm volatile("fence.acq_rel.sys;" ::: "memory");  // the CTA's peer stores are performed before it arrives
      else asm volatile("fence.acq_rel.gpu;" ::: "memory");                       // the CTA's vector writes (bar.sync: cumulative)
      ll_store_gpu(ll + 4 * blockIdx.x + 2 * lane, (unsigned long long)__double_as_longlong(lane == 0 ? c0 : c1), g32);
    }
    if (timed && lane == 0) ns_arrive = global_ns();
  }
  if (master) {
    // ---- master CTA: fetch every CTA's partials (the poll IS the arrival wait) ----
    double s0 = 0.0, s1 = 0.0;
    bool okp = true;
    const long long c0k = clock64();
    for (unsigned i = tid; i < G; i += NCT) {
      unsigned long long b0 = 0ull, b1 = 0ull;
      while (!(ll_load_gpu(ll + 4 * i, g32, b0) & ll_load_gpu(ll + 4 * i + 2, g32, b1)))
        if (clock64() - c0k > a.spin_limit) {
          okp = false;
          break;
        }
      s0 += __longlong_as_double((long long)b0);
      s1 += __longlong_as_double((long long)b1);
    }
    asm volatile("fence.acq_rel.gpu;" ::: "memory");
    s0 = warp_sum(s0);
    s1 = warp_sum(s1);
    okp = __all_sync(0xffffffffu, okp);
    if (lane == 0) {
      red[warp] = s0;
      red[16 + warp] = s1;
    }
    const bool ok_halo = red[35] != 0.0;  // written before the first CTA barrier of this call
    const bool timed_out = bar_or_consumers<NCT>(!okp);
    if (warp == 0) {
      s0 = (lane < NCW) ? red[lane] : 0.0;
      s1 = (lane < NCW) ? red[16 + lane] : 0.0;
      int ok = ok_halo && !timed_out;
      s0 = warp_sum(s0);
      s1 = warp_sum(s1);
      if (timed && lane == 0) {
        const unsigned long long now = global_ns();
        red64[FRED_T_LOCAL] += now - ns_arrive;
        ns_arrive = now;
      }
      const FusedPeer& P = a.peer;
      if (P.world > 1) {
        // lanes 0..15 carry `kind`, lanes 16..31 carry FB_RR when `both` (world <= 16)
        const int half = lane >> 4, pr = lane & 15;
        const int mykind = half ? FB_RR : kind;
        unsigned long long ep = 0;
        if (pr == 0 && (half == 0 || both)) ep = ++red64[FRED_EPOCH + mykind];
        ep = __shfl_sync(0xffffffffu, ep, lane & 16);
        const unsigned long long e32 = ep & 0xffffffffull;
        const double sv = half ? s1 : ((kind == FB_RR) ? s1 : s0);
        double rv = 0.0;
        if (pr < P.world && (half == 0 || both)) {
          const unsigned long long bits = (unsigned long long)__double_as_longlong(sv);
          if (fenced) asm volatile("fence.acq_rel.sys;" ::: "memory");  // after the acquired arrivals, before the words leave
          st_relaxed_sys_v2(&P.xchg[pr]->w[mykind][P.rank][0], ((bits & 0xffffffffull) << 32) | e32, (bits & 0xffffffff00000000ull) | e32);
          const unsigned long long* mine = &P.xchg[P.rank]->w[mykind][pr][0];
          unsigned long long w0v = 0ull, w1v = 0ull;
          const long long c1k = clock64();
          while (true) {
            ld_relaxed_sys_v2(mine, w0v, w1v);
            if ((w0v & 0xffffffffull) == e32 && (w1v & 0xffffffffull) == e32) break;
            if (clock64() - c1k > a.spin_limit) {
              ok = 0;
              break;
            }
          }
          if (fenced) asm volatile("fence.acq_rel.sys;" ::: "memory");
          rv = __longlong_as_double((long long)((w1v & 0xffffffff00000000ull) | (w0v >> 32)));
        }
        double tot = 0.0, tot2 = 0.0;
        for (int r = 0; r < P.world; ++r) {  // rank order: identical bits on every rank
          tot += __shfl_sync(0xffffffffu, rv, r);
          tot2 += __shfl_sync(0xffffffffu, rv, 16 + r);
        }
        if (kind == FB_RR) s1 = tot;
        else s0 = tot;
        if (both) s1 = tot2;
        if (timed && lane == 0) {
          red64[FRED_T_PEER] += global_ns() - ns_arrive;
          red64[FRED_N_BAR] += 1ull;
        }
      }
      ok = __all_sync(0xffffffffu, ok);
      if (lane < 3) {
        const unsigned long long bits = lane == 0 ? (unsigned long long)__double_as_longlong(s0)
                                       : lane == 1 ? (unsigned long long)__double_as_longlong(s1) : (ok ? ~0ull : 0ull);
        asm volatile("fence.acq_rel.gpu;" ::: "memory");  // everything the arrivals released is ordered before the words
        ll_store_gpu(&a.sync->rel[2 * lane], bits, g32);
      }
    }
  }
  if (warp == 0) {
    // every CTA: lanes 0..2 poll one pair each until both words carry this generation
    unsigned long long bits = 0ull;
    bool seen = true;
    if (lane < 3) {
      const long long c0k = clock64();
      while (!ll_load_gpu(&a.sync->rel[2 * lane], g32, bits))
        if (clock64() - c0k > 3 * a.spin_limit) {
          seen = false;
          break;
        }
      asm volatile("fence.acq_rel.gpu;" ::: "memory");
    }
    seen = __all_sync(0xffffffffu, seen);
    if (lane < 2) red[32 + lane] = __longlong_as_double((long long)bits);
    if (lane == 2) red[34] = (seen && bits != 0ull) ? 1.0 : 0.0;
  }
  bar_consumers<NCT>();
  t0 = red[32];
  t1 = red[33];
  const bool ok = red[34] != 0.0;
  return ok;
}

template <int CAP, int STAGES, typename VT, typename CT>
struct FusedSmem {
  using C = StreamCfg<CAP, VT, CT>;
  VT* vals;
  unsigned char* cols;
  int* rptr;
  int* first;
  int4* hdr;
  uint64_t *full, *empty;
};

// Value-less local grid barrier (consumer threads of every CTA).  Returns false on timeout.
// halo_epoch != 0 (row-partitioned, after t = G^T r): CTA 0 first waits until every neighbour whose rows this rank's halo
// holds has flagged that sweep (FusedXchg::hflag, raised by the neighbour's last boundary CTA -- early in ITS sweep, the
// boundary blocks go first), fences at system scope and only then arrives: the local release/acquire chain hands the
// halo data on to every CTA.  This replaces an all-to-all exchange by a (normally already satisfied) neighbour flag.
template <int NCW>
__device__ __forceinline__ bool fused_light_barrier(const FusedArgs& a, unsigned long long& lgen, double* red,
                                                    unsigned long long halo_epoch = 0ull) {
  constexpr int NCT = NCW * 32;
  lgen += 1;
  bar_consumers<NCT>();
  if (threadIdx.x == 0) {
    bool halo_ok = true;
    if (halo_epoch != 0ull && blockIdx.x == 0 && a.peer.n_recv > 0) halo_ok = fused_wait_halo(a, 0, halo_epoch);
    // arrivals and polls use different lines: the CTA whose arrival completes the count (the atomic's return value tells
    // it) publishes the generation; everybody else polls that word, undisturbed by the atomics
    bool seen = true;
    unsigned long long old = 0ull;
    if (halo_ok) asm volatile("atom.acq_rel.gpu.global.add.u64 %0, [%1], 1;" : "=l"(old) : "l"(&a.sync->light) : "memory");
    if (halo_ok && old + 1ull == lgen * (unsigned long long)gridDim.x) {
      st_release_gpu_u64(&a.sync->light_rel, lgen);
    } else {
      const long long w0 = clock64();
      while (ld_acquire_gpu_u64(&a.sync->light_rel) < lgen)
        if (clock64() - w0 > a.spin_limit) {
          seen = false;
          break;
        }
    }
    red[35] = seen ? 1.0 : 0.0;
  }
  bar_consumers<NCT>();
  return red[35] != 0.0;
}

// One matrix sweep by the consumer threads: blocks first, first + G, ... of M; `gv` is the gathered (extended) vector.
//   PH = 1 (G^T):  t = sum [/ d]
//   PH = 2 (G):    z = sum + eps r [/ d] ; acc += r z
//   PH = 3 (A):    q = sum ; acc += p q
// Called by every consumer thread of a CTA that owns boundary blocks, right after the last of them in a sweep of kind
// hk (0: t = G^T r, 1: z = G t, 2: q = A p): the CTA's peer stores are fenced at system scope, then counted; the CTA that
// completes the count raises this sweep's epoch in hflag[hk][my rank] of every neighbour it pushes to (one 8-byte
// NVLink store each).  This happens EARLY in the sweep -- the boundary blocks are dealt first -- so neither the fence nor
// the neighbours' wait for the flag sits on the critical path of the barrier that closes the sweep.
template <int NCW>
__device__ __forceinline__ void fused_boundary_done(const FusedArgs& a, const FusedMat& M, int hk, unsigned long long halo_epoch,
                                                    unsigned long long sweep_no) {
  bar_consumers<NCW * 32>();  // every consumer thread's stores of the boundary blocks are issued
  if (threadIdx.x == 0) {
    asm volatile("fence.acq_rel.sys;" ::: "memory");  // cumulative over the CTA (bar.sync): the peer stores are performed
    const unsigned long long n_bnd_ctas = (unsigned long long)min(M.nbnd, (int)gridDim.x);
    unsigned long long old;
    asm volatile("atom.acq_rel.gpu.global.add.u64 %0, [%1], 1;" : "=l"(old) : "l"(&a.sync->bdone[hk][0]) : "memory");
    if (old + 1ull == sweep_no * n_bnd_ctas) {  // the last boundary CTA of this sweep
      asm volatile("fence.acq_rel.sys;" ::: "memory");
      for (int s = 0; s < a.peer.n_push; ++s)
        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(&a.peer.xchg[a.peer.push_peer[s]]->hflag[hk][a.peer.rank]), "l"(halo_epoch) : "memory");
    }
  }
}

template <int PH, int CAP, int STAGES, int NCW, typename VT, typename CT>
__device__ __forceinline__ void fused_sweep(const FusedArgs& a, const FusedMat& M, const FusedSmem<CAP, STAGES, VT, CT>& sm,
                                            long long& item, const double* gv, double& acc, bool& pushed,
                                            unsigned long long halo_epoch = 0ull, unsigned long long sweep_no = 0ull) {
  using C = StreamCfg<CAP, VT, CT>;
  constexpr bool DELTA = C::DELTA;
  constexpr int NCT = NCW * 32;
  constexpr int CA = 16 / (int)sizeof(CT);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int G = gridDim.x, firstb = blockIdx.x;
  const int n_items = (firstb < M.nblk) ? (M.nblk - firstb + G - 1) / G : 0;
  const VT* Aval = reinterpret_cast<const VT*>(M.sval);
  const int off = a.off;
  const bool scaled = a.d != nullptr;
  const int np = a.peer.n_push;
  const double* own = (PH == 2) ? a.r : a.p;  // the owner-side operand of the dot (PH 2, 3)

  auto finish = [&](int i, double sum, double o0) {
    const int e = off + i;
    double out;
    double* const* dst;
    if (PH == 1) {
      out = scaled ? sum / a.d[i] : sum;
      a.t[e] = out;
      dst = a.peer.push_t;
    } else if (PH == 2) {
      double ev = __dmul_rn(a.eps, o0);
      if (scaled) ev = ev / a.d[i];
      out = __dadd_rn(sum, ev);
      a.z[e] = out;
      acc += o0 * out;
      dst = a.peer.push_z;
    } else {
      out = sum;
      a.q[e] = out;
      acc += o0 * out;
      dst = a.peer.push_q;
    }
    for (int s = 0; s < np; ++s)
      if (i >= a.peer.push_lo[s] && i < a.peer.push_hi[s]) {
        dst[s][i - a.peer.push_lo[s]] = out;
        pushed = true;
      }
  };

  // (row-partitioned) boundary blocks are dealt first: block index firstb + it * G < M.nbnd.  The CTA reports them
  // BND_LAG blocks later: by then its peer stores have long been acknowledged and the system-scope fence finds nothing in
  // flight (right behind the stores it costs an NVLink round trip with one thread of the CTA stalled)
  constexpr int BND_LAG = 2;
  const int n_bnd_items = (a.peer.world > 1 && firstb < M.nbnd) ? (M.nbnd - firstb + G - 1) / G : 0;
  for (int it = 0; it < n_items; ++it, ++item) {
    if (n_bnd_items > 0 && it == n_bnd_items + BND_LAG) fused_boundary_done<NCW>(a, M, PH - 1, halo_epoch, sweep_no);
    const int slot = (int)(item % STAGES);
    mbar_wait(sm.full + slot, (uint32_t)((item / STAGES) & 1));
    const int4 info = sm.hdr[slot];
    const int r0 = info.x, lo = info.y, r1 = info.z, hi = info.w;
    const int cnt = hi - lo;
    if (cnt > 0 && cnt <= CAP) {
      const VT* vals = sm.vals + slot * C::VALS + (lo & 3);
      const CT* cols = reinterpret_cast<const CT*>(sm.cols + slot * C::COLS_BYTES) + (lo & (CA - 1));
      const int* rptr = sm.rptr + slot * C::RPTR + (r0 & 3);
      const int* rfirst = sm.first + slot * C::RPTR + (r0 & 3);
      for (int rr = tid; rr < r1 - r0; rr += NCT) {
        const int row = r0 + rr;
        const double o0 = (PH == 1) ? 0.0 : ld_coh(own + off + row);
        int k = rptr[rr] - lo;
        const int e = rptr[rr + 1] - lo;
        int cprev = DELTA ? rfirst[rr] : 0;
        double sum = 0.0;
        // 16 entries (a mesh row) per pass: all gathers of a pass are in flight before the first add
        constexpr int U = 16;
#pragma unroll 1
        for (; k < e; k += U) {
          int c[U];
          double pr[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int raw = (int)cols[k + u];
            if (DELTA) {
              cprev += (k + u < e) ? raw : 0;
            } else {
              cprev = (k + u < e) ? raw : cprev;
            }
            c[u] = cprev;
          }
#pragma unroll
          for (int u = 0; u < U; ++u) pr[u] = __dmul_rn((k + u < e) ? (double)vals[k + u] : 0.0, gv[c[u]]);
#pragma unroll
          for (int u = 0; u < U; ++u) sum = __dadd_rn(sum, pr[u]);
        }
        finish(row, sum, o0);
      }
    } else {
      // a row longer than the stage capacity (or a block of empty rows): warp per row from global memory, products
      // added in entry order so these rows too carry scipy's bits
      for (int row = r0 + warp; row < r1; row += NCW) {
        const int s = __ldg(M.rowptr + row), e = __ldg(M.rowptr + row + 1);
        double sum = 0.0;
        for (int k0 = s; k0 < e; k0 += 32) {
          const int k = k0 + lane;
          const double pv = (k < e) ? __dmul_rn((double)__ldg(Aval + k), ld_coh(gv + __ldg(M.col + k))) : 0.0;
          const int m = min(32, e - k0);
          for (int j = 0; j < m; ++j) sum = __dadd_rn(sum, __shfl_sync(0xffffffffu, pv, j));
        }
        if (lane == 0) finish(row, sum, (PH == 1) ? 0.0 : ld_coh(own + off + row));
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(sm.empty + slot);
  }
  if (n_bnd_items > 0 && n_items <= n_bnd_items + BND_LAG) fused_boundary_done<NCW>(a, M, PH - 1, halo_epoch, sweep_no);
}

// MINB = resident CTAs per SM the register allocation is held to (4 -> 56 registers: the occupancy of the standalone
// streaming kernel, whose gathers need the warps; 3 -> 72 registers)
template <int CAP, int STAGES, int NCW, typename VT, typename CT, int MINB>
__global__ void __launch_bounds__((NCW + 1) * 32, MINB) pcg_fused_kernel(const __grid_constant__ FusedArgs a) {
  using C = StreamCfg<CAP, VT, CT>;
  constexpr bool DELTA = C::DELTA;
  constexpr int NCT = NCW * 32;
  constexpr int CA = 16 / (int)sizeof(CT);
  static_assert(NCW <= 16, "reduction scratch holds 16 warps");
  extern __shared__ __align__(128) unsigned char smem_raw[];
  unsigned char* sp = smem_raw;
  FusedSmem<CAP, STAGES, VT, CT> sm;
  sm.vals = reinterpret_cast<VT*>(sp);               sp += STAGES * C::VALS_BYTES;
  sm.cols = sp;                                      sp += STAGES * C::COLS_BYTES;
  sm.rptr = reinterpret_cast<int*>(sp);              sp += STAGES * C::RPTR * 4;
  sm.first = reinterpret_cast<int*>(sp);             sp += DELTA ? STAGES * C::RPTR * 4 : 0;
  sm.hdr = reinterpret_cast<int4*>(sp);              sp += STAGES * 16;
  sm.full = reinterpret_cast<uint64_t*>(sp);         sp += STAGES * 8;
  sm.empty = reinterpret_cast<uint64_t*>(sp);        sp += STAGES * 8;
  double* red = reinterpret_cast<double*>(sp);       sp += FRED_SLOTS * 8;
  volatile int* s_done = reinterpret_cast<volatile int*>(sp);

  const int tid = threadIdx.x;
  const int G = gridDim.x;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(sm.full + s, 1);
      mbar_init(sm.empty + s, NCW);
    }
    mbar_fence_init();
    *s_done = 0;
    unsigned long long* red64 = reinterpret_cast<unsigned long long*>(red);
    for (int k = 0; k < FB_KINDS; ++k) red64[FRED_EPOCH + k] = (blockIdx.x == 0 && a.peer.world > 1) ? a.peer.epoch[k] : 0ull;
    red64[FRED_T_LOCAL] = red64[FRED_T_PEER] = red64[FRED_N_BAR] = 0ull;
  }
  __syncthreads();

  if (tid >= NCT) {
    // ------------------------------------------------ producer warp ------------------------------------------------
    if (tid != NCT) return;
    const uint64_t policy = policy_evict_first();
    long long item = 0;
    const int firstb = blockIdx.x;
    bool stop = false;
    while (!stop) {
      bool any = false;
      for (int ph = 0; ph < 3 && !stop; ++ph) {
        const FusedMat& M = (ph == 0) ? a.GT : (ph == 1) ? a.G : a.A;
        const VT* Aval = reinterpret_cast<const VT*>(M.sval);
        const CT* Acol = reinterpret_cast<const CT*>(M.scol);
        const int n_items = (firstb < M.nblk) ? (M.nblk - firstb + G - 1) / G : 0;
        for (int j = 0; j < n_items; ++j, ++item) {
          any = true;
          const int slot = (int)(item % STAGES);
          if (item >= STAGES) {
            const uint32_t par = (uint32_t)(((item / STAGES) - 1) & 1);
            while (!mbar_try_wait(sm.empty + slot, par))
              if (*s_done) {
                stop = true;
                break;
              }
            if (stop) break;
          }
          const int4 cur = __ldg(M.blk + firstb + (long long)j * G);
          sm.hdr[slot] = cur;
          const int r0 = cur.x, lo = cur.y, r1 = cur.z, hi = cur.w;
          const int cnt = hi - lo;
          if (cnt > CAP || cnt <= 0) {
            mbar_arrive(sm.full + slot);
            continue;
          }
          const int v_lo = lo & ~3;
          const int vn = (hi - v_lo + 3) & ~3;
          const int c_lo = lo & ~(CA - 1);
          const int cn = (hi - c_lo + CA - 1) & ~(CA - 1);
          const int ra = r0 & ~3;
          const int nr4 = (r1 + 1 - ra + 3) & ~3;
          mbar_arrive_expect_tx(sm.full + slot, (uint32_t)(vn * (int)sizeof(VT) + cn * (int)sizeof(CT) + nr4 * 4 * (DELTA ? 2 : 1)));
          bulk_g2s(sm.vals + slot * C::VALS, Aval + v_lo, (uint32_t)(vn * sizeof(VT)), sm.full + slot, policy);
          bulk_g2s(sm.cols + slot * C::COLS_BYTES, Acol + c_lo, (uint32_t)(cn * sizeof(CT)), sm.full + slot, policy);
          bulk_g2s(sm.rptr + slot * C::RPTR, M.rowptr + ra, (uint32_t)nr4 * 4, sm.full + slot, policy);
          if (DELTA) bulk_g2s(sm.first + slot * C::RPTR, M.rowfirst + ra, (uint32_t)nr4 * 4, sm.full + slot, policy);
        }
      }
      if (!any) {  // a CTA without row blocks only takes part in the barriers and the vector phases
        while (!*s_done) __nanosleep(200);
        stop = true;
      }
    }
    // copies issued ahead of a solve that ended must land before the CTA's shared memory goes away
    for (long long k = (item > STAGES ? item - STAGES : 0); k < item; ++k)
      mbar_wait(sm.full + (int)(k % STAGES), (uint32_t)((k / STAGES) & 1));
    return;
  }

  // -------------------------------------------------- consumer warps --------------------------------------------------
  const int n = a.n, off = a.off;
  const int n_halo = off + a.hi_halo;
  const bool lead = (blockIdx.x == 0 && tid == 0);
  const long long gtid = (long long)blockIdx.x * NCT + tid, gstride = (long long)G * NCT;
  PcgScalars* sc = a.sc;
  const double rtol = sc->rtol;
  const int max_iter = sc->max_iter;
  unsigned long long gen = 0, lgen = 0;
  long long item = 0;
  double t0 = 0.0, t1 = 0.0;
  bool ok = true;

  // (row-partitioned) nobody stores into a peer before every rank is inside this solve
  if (a.peer.world > 1) ok = fused_barrier<NCW>(a, gen, FB_READY, 0.0, 0.0, false, red, t0, t1);

  // ---- init: r = b (owned; boundary entries also into the neighbours' halos), x = 0, ||b||^2 ----
  double acc = 0.0;
  bool pushed = false;
  if (ok) {
    for (long long i = gtid; i < n; i += gstride) {
      const double v = a.b[i];
      a.r[off + i] = v;
      a.x[i] = 0.0;
      acc += v * v;
      for (int s = 0; s < a.peer.n_push; ++s)
        if (i >= a.peer.push_lo[s] && i < a.peer.push_hi[s]) {
          a.peer.push_r[s][i - a.peer.push_lo[s]] = v;
          pushed = true;
        }
    }
    ok = fused_barrier<NCW>(a, gen, FB_INIT, acc, 0.0, pushed, red, t0, t1);
  }
  const double bb = t0;
  const double bnorm = sqrt(bb);
  const double atol = fmax(0.0, rtol * bnorm);
  double rr = bb, rho = 0.0, rho_prev = 0.0, pq = 0.0, alpha = 0.0, beta = 0.0;
  int iter = 0, done = 0, converged = 0, breakdown = 0;
  if (!ok) {
    done = 1;
    breakdown = 3;
  } else {
    if (lead) a.hist[0] = bnorm;
    if (bnorm == 0.0) {
      done = 1;
      converged = 1;
    } else if (!(bnorm == bnorm)) {
      done = 1;
      breakdown = 1;
    }
  }
  bool pending = false;   // an x / r update was made: its ||r||^2 (acc_rr) is all-reduced with barrier B1
  double acc_rr = 0.0;
  const bool prof = lead && a.prof != nullptr;
  unsigned long long ns_prev = prof ? global_ns() : 0ull, ns_p[5] = {0ull, 0ull, 0ull, 0ull, 0ull};
  auto stamp = [&](int k) {
    if (prof) {
      const unsigned long long now = global_ns();
      ns_p[k] += now - ns_prev;
      ns_prev = now;
    }
  };

  // Row-partitioned: the hand-over of t between P1 and P2 needs the local grid and the NEIGHBOURS' boundary rows only,
  // so barrier B1 is a local barrier + neighbour flags, and ||r||^2 of the previous update travels with rho in B2 (the
  // P2 sweep that runs ahead of a converged test only writes z).
  // the boundaries of z and q are flagged the same way, so the scalar all-reduces B2 / B3 carry no system-scope fence
  const bool multi = a.peer.world > 1;
  const unsigned long long halo_base[3] = {multi ? a.peer.epoch[FB_HALO] : 0ull, multi ? a.peer.epoch[FB_HALO + 1] : 0ull,
                                           multi ? a.peer.epoch[FB_HALO + 2] : 0ull};
  unsigned long long n_p1 = 0, n_p2 = 0, n_p3 = 0;

  while (!done) {
    // ---- P1: t = G^T r ----
    acc = 0.0;
    pushed = false;
    n_p1 += 1;
    fused_sweep<1, CAP, STAGES, NCW, VT, CT>(a, a.GT, sm, item, a.r, acc, pushed, halo_base[0] + n_p1, n_p1);
    ok = fused_light_barrier<NCW>(a, lgen, red, multi ? halo_base[0] + n_p1 : 0ull);
    stamp(0);
    if (!ok) {
      breakdown = 3;
      break;
    }
    // ---- P2: z = G t + eps r ; rho ----
    acc = 0.0;
    pushed = false;
    n_p2 += 1;
    fused_sweep<2, CAP, STAGES, NCW, VT, CT>(a, a.G, sm, item, a.t, acc, pushed, halo_base[1] + n_p2, n_p2);
    ok = fused_barrier<NCW>(a, gen, FB_RHO, acc, pending ? acc_rr : 0.0, false, red, t0, t1, true, 1, halo_base[1] + n_p2);
    stamp(1);
    if (!ok) {
      breakdown = 3;
      break;
    }
    if (pending) {  // the update is complete: count it, test it (scipy: callback, then the test at the loop top)
      if (!(t1 == t1)) {
        breakdown = 1;
        break;
      }
      rr = t1;
      rho_prev = rho;
      iter += 1;
      const double nrm = sqrt(t1);
      if (lead) a.hist[iter] = nrm;
      if (nrm < atol) {
        converged = 1;
        break;
      }
      if (iter >= max_iter) break;
    }
    rho = t0;
    const bool first = (iter == 0);
    beta = first ? 0.0 : rho / rho_prev;
    // ---- V2: p = z + beta p (owned rows and halo; scipy: p *= beta; p += z) ----
    {
      // the extended vectors start 16-byte aligned: two entries per access, L2-coherent (.cg) loads and stores
      const long long n_ext = (long long)n_halo + n;
      const long long n2 = n_ext >> 1;
      double2* p2 = reinterpret_cast<double2*>(a.p);
      const double2* z2 = reinterpret_cast<const double2*>(a.z);
      if (first) {
#pragma unroll 2
        for (long long j = gtid; j < n2; j += gstride) __stcg(p2 + j, __ldcg(z2 + j));
        if (gtid == 0 && (n_ext & 1)) a.p[n_ext - 1] = __ldcg(a.z + n_ext - 1);
      } else {
#pragma unroll 2
        for (long long j = gtid; j < n2; j += gstride) {
          const double2 pv = __ldcg(p2 + j), zv = __ldcg(z2 + j);
          __stcg(p2 + j, make_double2(__dadd_rn(__dmul_rn(pv.x, beta), zv.x), __dadd_rn(__dmul_rn(pv.y, beta), zv.y)));
        }
        if (gtid == 0 && (n_ext & 1)) a.p[n_ext - 1] = __dadd_rn(__dmul_rn(__ldcg(a.p + n_ext - 1), beta), __ldcg(a.z + n_ext - 1));
      }
    }
    ok = fused_light_barrier<NCW>(a, lgen, red);
    stamp(2);
    if (!ok) {
      breakdown = 3;
      break;
    }
    // ---- P3: q = A p ; p.q ----
    acc = 0.0;
    pushed = false;
    n_p3 += 1;
    fused_sweep<3, CAP, STAGES, NCW, VT, CT>(a, a.A, sm, item, a.p, acc, pushed, halo_base[2] + n_p3, n_p3);
    ok = fused_barrier<NCW>(a, gen, FB_PQ, acc, 0.0, false, red, t0, t1, false, 2, halo_base[2] + n_p3);
    stamp(3);
    if (!ok) {
      breakdown = 3;
      break;
    }
    pq = t0;
    if (!(pq > 0.0) || !(rho == rho)) {  // p.q <= 0 or NaN: A or M not SPD
      breakdown = 1;
      alpha = 0.0;
      break;
    }
    alpha = rho / pq;
    // ---- V1: x += alpha p ; r -= alpha q (owned rows, then the halo of r) ; ||r||^2 over the owned rows ----
    acc_rr = 0.0;
    {
      // the owned slice starts at an even extended index (16-byte aligned); x is the caller's array and may not be
      const bool x16 = (reinterpret_cast<unsigned long long>(a.x) & 15ull) == 0ull;
      const long long n2 = x16 ? (long long)(n >> 1) : 0;
      const double2* p2 = reinterpret_cast<const double2*>(a.p + off);
      const double2* q2 = reinterpret_cast<const double2*>(a.q + off);
      double2* r2 = reinterpret_cast<double2*>(a.r + off);
      double2* x2 = reinterpret_cast<double2*>(a.x);
#pragma unroll 2
      for (long long j = gtid; j < n2; j += gstride) {
        const double2 pv = __ldcg(p2 + j), qv = __ldcg(q2 + j), rv = __ldcg(r2 + j), xv = __ldcg(x2 + j);
        __stcg(x2 + j, make_double2(__dadd_rn(xv.x, __dmul_rn(alpha, pv.x)), __dadd_rn(xv.y, __dmul_rn(alpha, pv.y))));
        const double r0 = __dsub_rn(rv.x, __dmul_rn(alpha, qv.x)), r1 = __dsub_rn(rv.y, __dmul_rn(alpha, qv.y));
        __stcg(r2 + j, make_double2(r0, r1));
        acc_rr += r0 * r0;
        acc_rr += r1 * r1;
      }
      for (long long i = 2 * n2 + gtid; i < n; i += gstride) {  // the odd last entry, or everything when x is unaligned
        const long long e = off + i;
        a.x[i] = __dadd_rn(__ldcg(a.x + i), __dmul_rn(alpha, __ldcg(a.p + e)));
        const double rv = __dsub_rn(__ldcg(a.r + e), __dmul_rn(alpha, __ldcg(a.q + e)));
        a.r[e] = rv;
        acc_rr += rv * rv;
      }
    }
    for (long long h = gtid; h < n_halo; h += gstride) {
      const long long e = h < off ? h : (long long)n + h;
      a.r[e] = __dsub_rn(__ldcg(a.r + e), __dmul_rn(alpha, __ldcg(a.q + e)));
    }
    pending = true;
    ok = fused_light_barrier<NCW>(a, lgen, red);
    stamp(4);
    if (!ok) {
      breakdown = 3;
      break;
    }
  }

  if (tid == 0) *s_done = 1;  // the producer stops prefetching and drains
  if (prof) {
    for (int k = 0; k < 5; ++k) a.prof[k] = ns_p[k];
    a.prof[5] = (unsigned long long)iter;
    const unsigned long long* red64 = reinterpret_cast<const unsigned long long*>(red);
    a.prof[6] = red64[FRED_T_LOCAL];
    a.prof[7] = red64[FRED_T_PEER];
    a.prof[8] = red64[FRED_N_BAR];
  }
  if (lead && a.peer.world > 1) {  // the epoch counters survive the solve (monotonic across solves and relayouts)
    const unsigned long long* red64 = reinterpret_cast<const unsigned long long*>(red);
    for (int k = 0; k < FB_KINDS; ++k) a.peer.epoch[k] = red64[FRED_EPOCH + k];
    a.peer.epoch[FB_HALO] = halo_base[0] + n_p1;
    a.peer.epoch[FB_HALO + 1] = halo_base[1] + n_p2;
    a.peer.epoch[FB_HALO + 2] = halo_base[2] + n_p3;
  }
  if (lead) {
    sc->bnorm = bnorm;
    sc->atol = atol;
    sc->rr = rr;
    sc->rho = rho;
    sc->rho_prev = rho_prev;
    sc->pq = pq;
    sc->alpha = alpha;
    sc->beta = beta;
    sc->iter = iter;
    sc->done = 1;
    sc->converged = converged;
    sc->breakdown = breakdown;
  }
}

// host side (pcg_fused.cu)
struct FusedPlan {
  int minb = 4;     // kernel variant: resident CTAs per SM
  int which = -1;   // partition index (1: 2048-entry stages, 2: 4096)
  int fmt = 0;      // bit 0: fp32 values, bit 1: 16-bit column steps
  int grid = 0;     // co-resident CTAs
  int threads = 0;
};
// Builds the lazily created pieces of the three matrices and decides whether the fused kernel applies
// (plan->which < 0: not eligible -- the caller falls back to the launch-per-kernel drivers).  Synchronises `st`.
int fused_plan(b200spai_csr_s* A, b200spai_csr_s* G, b200spai_csr_s* GT, cudaStream_t st, FusedPlan* plan);
FusedMat fused_mat(const b200spai_csr_s* M, const FusedPlan& plan);
int fused_launch(const FusedArgs& args, const FusedPlan& plan, cudaStream_t st);

}  // namespace b200spai
