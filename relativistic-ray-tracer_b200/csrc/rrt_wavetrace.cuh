// rrt_wavetrace.cuh — the trace kernel: BVHAccel::intersect (reference src/bvh.cpp:99-138) for a queue of rays.
//
// Persistent warps.  Rays need anywhere from 1 to 63 chords, so a warp that waited for its slowest ray would idle
// most lanes (measured on the first version: 9-12 active threads per instruction).  Here every LANE owns one ray
// and replaces it the moment it finishes: the warp reserves RRT_POOL consecutive queue entries with one atomic,
// hands them to free lanes with ballot + popc, and only goes back to the global cursor when its pool is empty.
// One loop trip = at most one chord per lane: geodesic advance, horizon test, BVH traversal of that chord.
//
// IO policies say where rays come from and where results go (extension queue, shadow queue, parity hook), so the
// three kernels run the same loop.
#pragma once
#include "rrt_trace.cuh"

#ifndef RRT_AB
#define RRT_AB 0   /* 1 = also compile the measured-slower A/B variants: the vote-scheduled loop (V = 1) and 4-wide nodes (make EXTRA=-DRRT_AB=1) */
#endif
#ifndef RRT_BLOCK
#define RRT_BLOCK 128
#endif
#ifndef RRT_POOL
#define RRT_POOL 32   /* queue entries a warp reserves per atomic.  Entries a warp has reserved but not started are invisible to the other warps: at the
                         end of a launch they are worked off lane by lane while the rest of the GPU idles.  128 -> 32 on B200 (profiles/r02_c_*): C4 on
                         8 GPUs 513 -> 477 ms per frame (small launches: 3 rays per lane), single GPU bunny 207.1 -> 204.5, C2 33.5 -> 32.5, C5 53.9 -> 51.4 */
#endif
#ifndef RRT_MINB
#define RRT_MINB 7     /* resident blocks per SM the binary-node trace kernels are compiled for (72 registers).  6 until the chord moved to shared
                          memory (RRT_SEG_SMEM) and the tree-top record brought spills back at 80 registers: 6 -> 7 with the chord in shared memory:
                          bunny in-box 174.4 -> 171.5 ms, C5 50.8 -> 50.0, bunny 1080p 77.8 -> 77.1 (profiles/r02_k_ab_smem_minb.jsonl) */
#endif
#ifndef RRT_COOP_LEAVES
#define RRT_COOP_LEAVES 0   /* 1 = any-hit leaf tests deferred and shared across the warp (rrt_trace.cuh coop_*): measured SLOWER, C2 41.4 vs 37.1 ms */
#endif
#ifndef RRT_SEG_SMEM
#define RRT_SEG_SMEM 1   /* the current chord (o, d, len) of the hole-on FP64 loop lives in shared memory between steps (7 KB per block): it is read by the
                            stepper, the box context and the leaf tests only, and holds no registers while nodes are visited.  Alone (6 blocks per SM) it
                            changes nothing; it is what lets a 7th block in without spills in the shadow kernel (RRT_MINB) */
#endif
#ifndef RRT_MINB_FP32
#define RRT_MINB_FP32 5   /* resident blocks per SM for the FP32-stepper variants (STEP = 1): 4 / 5 / 6 / 7 / 8 measured, see DESIGN.md 6.2 */
#endif
#define RRT_MINBLOCKS(V, STEP) ((STEP) == 1 ? RRT_MINB_FP32 : (V) >= 4 ? RRT_MINB_WIDE : RRT_MINB)
#ifndef RRT_MINB_WIDE
#define RRT_MINB_WIDE 6   /* the same for the wide-node variants */
#endif

__device__ __forceinline__ int wt_lane() { return threadIdx.x & 31; }

/* hit geometry as Triangle::intersect / Sphere::intersect store it (reference src/static_scene/triangle.cpp:46-52,
 * sphere.cpp:44-50): hit_p = o + d t, w_out = -d, n = barycentric blend of the vertex normals (NOT normalised) or
 * the unit sphere normal */
__device__ __forceinline__ void fill_geo(const DevScene& sc, const Seg& s, const Best& best, Hit& hit) {
  hit.p = vadd(s.o, vmul(s.d, best.t));
  hit.w_out = vneg(s.d);
  uint32_t idx = sc.prim_index[best.prim];
  if (best.kind == 0) {
    const double* N = sc.tri_nrm + 9 * (size_t)idx;
    double b0 = 1 - best.b1 - best.b2;
    hit.n = vadd(vadd(smul(b0, D3(N[0], N[1], N[2])), smul(best.b1, D3(N[3], N[4], N[5]))), smul(best.b2, D3(N[6], N[7], N[8])));
  } else {
    const double* c = sc.sph + 4 * (size_t)idx;
    hit.n = vunit(vsub(vadd(s.o, vmul(s.d, best.t)), D3(c[0], c[1], c[2])));
  }
}

struct ClosestIO {   // extension queue -> hit records for the shade kernel
  const double* ray; size_t cap; int* h_prim; double* h_geo;
  static constexpr bool GEO = true;
  __device__ __forceinline__ void load(int i, d3& o, d3& d, double& tmin, double& tmax) const {
    o = D3(ray[i], ray[cap + i], ray[2 * cap + i]); d = D3(ray[3 * cap + i], ray[4 * cap + i], ray[5 * cap + i]);
    tmin = ray[6 * cap + i]; tmax = ray[7 * cap + i];
  }
  __device__ __forceinline__ void store(int i, const Hit& h) const {
    h_prim[i] = h.status == 1 ? h.prim : (h.status == 2 ? -2 : -1);
    if (h.status == 1) {
      h_geo[i] = h.p.x; h_geo[cap + i] = h.p.y; h_geo[2 * cap + i] = h.p.z;
      h_geo[3 * cap + i] = h.w_out.x; h_geo[4 * cap + i] = h.w_out.y; h_geo[5 * cap + i] = h.w_out.z;
      h_geo[6 * cap + i] = h.n.x; h_geo[7 * cap + i] = h.n.y; h_geo[8 * cap + i] = h.n.z;
    }
  }
};

struct ShadowIO {    // shadow queue -> verdict in the contribution table
  const double* ray; size_t cap; const uint32_t* cidx; int* res;
  const uint32_t* perm;   // launch order -> queue position (RRT_OPT_SORT_SHADOW, rrt_kernels.cuh k_bin_*), or nullptr: ray i is the i-th of the queue
  static constexpr bool GEO = false;
  __device__ __forceinline__ void load(int i, d3& o, d3& d, double& tmin, double& tmax) const {
    const size_t q = perm ? (size_t)perm[i] : (size_t)i;
    o = D3(ray[q], ray[cap + q], ray[2 * cap + q]); d = D3(ray[3 * cap + q], ray[4 * cap + q], ray[5 * cap + q]);
    tmin = 0.0; tmax = ray[6 * cap + q];
  }
  __device__ __forceinline__ void store(int i, const Hit& h) const { res[cidx[perm ? (size_t)perm[i] : (size_t)i]] = h.status == 1 ? h.prim : -1; }
};

template <bool ANY>
struct HookIO {      // rrt_trace_rays: host-layout arrays in, rrt_hit out
  const double* o; const double* d; const double* tmin; const double* tmax; rrt_hit* out;
  static constexpr bool GEO = !ANY;
  __device__ __forceinline__ void load(int i, d3& ro, d3& rd, double& mn, double& mx) const {
    ro = D3(o[3 * (size_t)i], o[3 * (size_t)i + 1], o[3 * (size_t)i + 2]); rd = D3(d[3 * (size_t)i], d[3 * (size_t)i + 1], d[3 * (size_t)i + 2]);
    mn = tmin ? tmin[i] : 0.0; mx = tmax ? tmax[i] : (double)INFINITY;
  }
  __device__ __forceinline__ void store(int i, const Hit& h) const {
    rrt_hit r;
    bool geo = h.status == 1 && !ANY;
    r.status = h.status; r.prim = geo ? h.prim : -1; r.chord = h.chord; r.steps = h.steps;
    r.t = h.status == 1 ? h.t : 0.0;
    r.hit_p[0] = geo ? h.p.x : 0; r.hit_p[1] = geo ? h.p.y : 0; r.hit_p[2] = geo ? h.p.z : 0;
    r.w_out[0] = geo ? h.w_out.x : 0; r.w_out[1] = geo ? h.w_out.y : 0; r.w_out[2] = geo ? h.w_out.z : 0;
    r.n[0] = geo ? h.n.x : 0; r.n[1] = geo ? h.n.y : 0; r.n[2] = geo ? h.n.z : 0;
    out[i] = r;
  }
};

__device__ __forceinline__ void wt_flush_counters(const Counters& c, Counters* g) {
  unsigned long long v[5] = {c.rays, c.steps, c.nodes, c.tris, c.sph};
  unsigned long long* dst = (unsigned long long*)g;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    unsigned long long x = v[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (wt_lane() == 0 && x) atomicAdd(dst + k, x);
  }
}

/* FP32, conservative: true when every point within `len` of p is outside the (outward-rounded) root box.  |p - fl(p)| <=
 * sqrt(3) 2^-24 max|p_i| < 1.04e-7 m enters the reach L; the float evaluation of the box distance is good to a few 2^-24
 * relative, covered by the 1e-5 margins.  NaN / inf anywhere compares false: the chord takes the traversal path. */
__device__ __forceinline__ bool chord_outside_root(const DevScene& sc, const d3& p, double len) {
  const float px = (float)p.x, py = (float)p.y, pz = (float)p.z;
  const float m = fmaxf(fmaxf(fabsf(px), fabsf(py)), fabsf(pz));
  const float ex = fmaxf(fmaxf(sc.root_lo[0] - px, px - sc.root_hi[0]), 0.f);
  const float ey = fmaxf(fmaxf(sc.root_lo[1] - py, py - sc.root_hi[1]), 0.f);
  const float ez = fmaxf(fmaxf(sc.root_lo[2] - pz, pz - sc.root_hi[2]), 0.f);
  const float d2 = ex * ex + ey * ey + ez * ez;
  const float L = __double2float_ru(len) * 1.00001f + m * 2.4e-7f;
  return d2 * 0.99999f > L * L;
}


/* ---------------------------------------------------------------- escape hand-off (round 2; VERDICT r01 item 3: "two-phase trace for rays outside the scene box")
 * A ray that has left the scene box keeps being integrated for all of its 63 chords (the reference never stops), and none of
 * those chords can hit anything while the box is out of their reach.  Carrying such a ray in the trace kernel wastes a lane of a warp whose
 * other lanes traverse the BVH, in an 80-register kernel that cannot reach the stepper's own throughput (28 vs 47 G chords/s).
 * So the trace kernel PARKS a ray at the first chord that starts outside the root box and is shorter than its distance to it
 * (chord_outside_root): (o, d, len, j, queue index) go to a park queue and the lane takes a fresh ray.  k_escape — stepper, horizon test and the
 * conservative root-box test only, no BVH, no stack, 8+ blocks per SM — then carries the parked rays on: captured -> result; out of chords ->
 * miss; a chord that passes the SAME root-box slab test traverse() starts with -> re-entry queue, finished by a second, small launch of the trace
 * kernel that starts with that chord's traversal (no parking there).  Every chord, horizon test and primitive test is the same arithmetic in the
 * same order; only chords that provably reach no primitive skip their (empty) traversal: results are bit-identical.
 * MEASURED SLOWER on B200 (profiles/r02_d_ab_escape.jsonl, ms per frame, off -> on): bunny 1080p 16 spp 83.9 -> 86.0, C5 51.0 -> 51.4, C2 32.2 -> 34.6,
 * bunny 480x360 64 spp 200.9 -> 212.1, C4 3311 -> 3422.  Escaped rays already run at 34 G chords/s in the trace kernel (2 streams, 32-entry pools) against
 * 48.5 G for stepper + horizon test ALONE from registers (tools/ilp_probe.cu): the lean kernel still pays the clear test, the queue round trip (128 B per
 * ray) and two more launch tails per trace phase, and the trace kernel pays the clear test on every chord.  The FP64 stepper, not the traversal state that
 * travels with it, bounds these rays.  Kept behind RRT_AB (RRT_OPT_ESCAPE = 1 in A/B builds). */
struct ParkQ {
  double* st;    /* 7 x cap: o.xyz d.xyz len of the last chord made */
  int* meta;     /* 2 x cap: caller's queue index, chords done so far (j) */
  int* cnt;      /* [0] entries written, [1] consumer cursor */
  size_t cap;
};

/* ---------------------------------------------------------------- variant 0 (default): chord-lockstep loop */
/* Escape warp (RRT_AB = 1 and RRT_ESCAPE_WARP = 1; hole on, FP64 stepper).  MEASURED SLOWER, not in the default build, kept as the record of the experiment.
 * In a launch of in-box rays a third of all chords belong to rays that have left the scene box through the open front and still walk their 63
 * chords (the reference never stops): they need no traversal, but each of them holds a lane of a warp whose other lanes do — the tree-top record
 * is visited at 20 of 32 lanes (profiles/r02_s_attribution_c4_shadow_final.txt).  With the flag on, warp 0 of every block takes such rays over: a
 * lane of the other warps whose chord missed the root box on two chords in a row hands its ray (chord, index, chord count: 64 B) to a 32-entry ring
 * in shared memory and takes a fresh one; warp 0 refills its lanes from that ring first and from the launch's queue otherwise.  Who executes a
 * chord changes, nothing else: frames are bit-identical (same SHA-256 on bunny_box, C2, C4).  A full ring leaves the ray where it is, and once the
 * launch's queue is exhausted nobody hands over any more.  Time, B200, same box, best of 3: bunny_box 165.3 -> 180.7 ms, C2 32.4 -> 35.7 ms,
 * C4 2684 -> 2925 ms (profiles/r02_u_escape_warp.txt): the traversal of a warp costs its slowest lane, so filling the idle lanes with in-box
 * rays lengthens every trip of three warps in four by more than the trips of the fourth get shorter. */
#ifndef RRT_ESCAPE_WARP
#define RRT_ESCAPE_WARP 0
#endif
#define RRT_EQ_CAP 32
struct EscQ {
  double st[7][RRT_EQ_CAP];
  int idx[RRT_EQ_CAP], j[RRT_EQ_CAP], flag[RRT_EQ_CAP];   /* flag: 1 = entry published */
  int head, tail, count, producers;                      /* head: consumer only; count >= tail - head at all times; producers: warps that may still publish */
};

/* PARK: 0 = every ray runs to its end here; 1 = escaping rays are parked in `pk` (hole on, FP64 stepper); 2 = the queue IS `pk` (re-entries from
 * k_escape: state + a chord whose traversal is pending), no parking */
template <bool ANY, bool COUNT, int W, int STEP, class IO, int PARK = 0>
__device__ __forceinline__ void trace_persistent(const DevScene& sc, const DevHole& bh, const IO& io, const int n, int* cursor, Counters* gcnt,
                                                 const ParkQ pk = ParkQ()) {
  __shared__ int stack_mem[RRT_STACK * RRT_BLOCK];
  __shared__ CoopScratch coop_mem[ANY ? RRT_BLOCK / 32 : 1];
  constexpr bool SMEM = RRT_SEG_SMEM && STEP == 0 && PARK == 0 && !RRT_COOP_LEAVES;   /* the chord lives in shared memory, not in registers */
  __shared__ double seg_mem[SMEM ? 7 * RRT_SEG_STRIDE : 1];
  double* sm = seg_mem + (SMEM ? threadIdx.x : 0);
  int* stack = stack_mem + threadIdx.x;
  constexpr bool EW = RRT_AB && RRT_ESCAPE_WARP && SMEM && RRT_BLOCK >= 64;   /* A/B builds only: make EXTRA="-DRRT_AB=1 -DRRT_ESCAPE_WARP=1" */
  __shared__ EscQ eq_mem[1];   /* unreferenced, hence dropped, when EW is false */
  EscQ& eq = eq_mem[0];
  const bool esc_warp = EW && (threadIdx.x >> 5) == 0;
  int out_streak = 0;                /* consecutive chords that missed the root box */
  if constexpr (EW) {
    if (threadIdx.x < RRT_EQ_CAP) eq.flag[threadIdx.x] = 0;
    if (threadIdx.x == 0) { eq.head = 0; eq.tail = 0; eq.count = 0; eq.producers = RRT_BLOCK / 32 - 1; }
    __syncthreads();
  }
  constexpr bool COOP = ANY && RRT_COOP_LEAVES;
  CoopScratch* coop = &coop_mem[COOP ? threadIdx.x >> 5 : 0];
  if (COOP) coop_init(coop);
  Counters cnt = {0, 0, 0, 0, 0, 0};
  constexpr bool flat = STEP == 2;   /* STEP: 0 = hole on, FP64 stepper; 1 = hole on, FP32 stepper; 2 = flat spacetime (the host dispatches on the mode:
                                        a compile-time switch keeps the flat path's tmin / tmax and the hole path's chord state out of each other's registers) */
  const unsigned lt_mask = (1u << wt_lane()) - 1u;
  int pool_next = 0, pool_end = 0;   /* warp-uniform */
  bool more = true;                  /* warp-uniform: the global queue may still have entries */
  bool has = false, skip = false;    /* skip (PARK 2): the loaded chord is already made and horizon-tested, its traversal comes first */
  int idx = 0, j = 0;
  FastRay fr; fr.xx = fr.xy = fr.xz = fr.yx = fr.yy = fr.yz = fr.u = fr.up = 0.f;
  float e0x = 0.f, e0y = 0.f, e0z = 0.f;   /* STEP 1: (double origin) - (float origin of the first chord) */
  d3 o = D3(0, 0, 0), d = D3(0, 0, 0);
  double len = 0.0, tmin = 0.0, tmax = 0.0;
  while (true) {
    /* ---- hand queue entries to the lanes that have none */
    unsigned need = __ballot_sync(0xffffffffu, !has);
    if constexpr (EW) {
      if (esc_warp && need) {   /* the ring first: rays the other warps of the block handed over */
        int h0 = 0, k = 0;
        if (wt_lane() == 0) {
          h0 = eq.head;
          const int want = __popc(need);
          while (k < want && ((volatile int*)eq.flag)[(h0 + k) % RRT_EQ_CAP] == 1) ++k;
        }
        h0 = __shfl_sync(0xffffffffu, h0, 0); k = __shfl_sync(0xffffffffu, k, 0);
        if (k > 0) {
          __threadfence_block();
          const int rank = __popc(need & lt_mask);
          if (!has && rank < k) {
            const int q = (h0 + rank) % RRT_EQ_CAP;
#pragma unroll
            for (int c = 0; c < 7; ++c) sm[c * RRT_SEG_STRIDE] = ((volatile double*)eq.st[c])[q];
            idx = ((volatile int*)eq.idx)[q]; j = ((volatile int*)eq.j)[q];
            has = true; out_streak = 2;
          }
          __syncwarp();
          if (wt_lane() == 0) {
            for (int r = 0; r < k; ++r) ((volatile int*)eq.flag)[(h0 + r) % RRT_EQ_CAP] = 0;
            eq.head = h0 + k;
            __threadfence_block();
            atomicSub(&eq.count, k);
          }
          need = __ballot_sync(0xffffffffu, !has);
        }
      }
    }
    while (need) {
      if (pool_next >= pool_end) {
        if (!more) break;
        int base = 0;
        if (wt_lane() == 0) base = atomicAdd(cursor, RRT_POOL);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) { more = false; break; }
        pool_next = base; pool_end = min(base + RRT_POOL, n);
      }
      const int avail = pool_end - pool_next;
      const int rank = __popc(need & lt_mask);
      if (!has && rank < avail) {
        if constexpr (PARK == 2) {
          const size_t q = (size_t)(pool_next + rank);
          o = D3(pk.st[q], pk.st[pk.cap + q], pk.st[2 * pk.cap + q]); d = D3(pk.st[3 * pk.cap + q], pk.st[4 * pk.cap + q], pk.st[5 * pk.cap + q]);
          len = pk.st[6 * pk.cap + q]; idx = pk.meta[q]; j = pk.meta[pk.cap + q];
          has = true; skip = true; tmin = 0.0; tmax = len;
        } else {
        idx = pool_next + rank;
        io.load(idx, o, d, tmin, tmax);
        has = true; j = 0; len = 0.0; out_streak = 0;
        if constexpr (SMEM) {
          sm[0] = o.x; sm[RRT_SEG_STRIDE] = o.y; sm[2 * RRT_SEG_STRIDE] = o.z; sm[3 * RRT_SEG_STRIDE] = d.x; sm[4 * RRT_SEG_STRIDE] = d.y; sm[5 * RRT_SEG_STRIDE] = d.z;
          sm[6 * RRT_SEG_STRIDE] = 0.0;
        }
        }
        if (STEP == 1) {
          fast_init(bh, o, d, fr);
          const float rho = fast_rcp(fr.u);   /* the float origin fast_step is about to form, to the bit */
          e0x = (float)(o.x - (double)__fmaf_rn(rho, fr.xx, bh.f_cx)); e0y = (float)(o.y - (double)__fmaf_rn(rho, fr.xy, bh.f_cy));
          e0z = (float)(o.z - (double)__fmaf_rn(rho, fr.xz, bh.f_cz));
        }
        if (COUNT && PARK != 2) cnt.rays++;
      }
      pool_next += min(__popc(need), avail);
      need = __ballot_sync(0xffffffffu, !has);
    }
    if (!__any_sync(0xffffffffu, has)) {
      if constexpr (EW) {
        if (esc_warp) {   /* nothing to do right now: done only when no warp of the block can publish any more and the ring is empty */
          int fin = 0;
          if (wt_lane() == 0) fin = (!more && *(volatile int*)&eq.producers == 0 && *(volatile int*)&eq.count == 0) ? 1 : 0;
          fin = __shfl_sync(0xffffffffu, fin, 0);
          if (fin) break;
          __nanosleep(200);
          continue;
        }
      }
      break;
    }
    /* ---- one chord for every lane that holds a ray */
    Hit h; h.status = -1;
    Best best;
    bool found;
    if constexpr (STEP == 1) {
      SegF sf; sf.ox = sf.oy = sf.oz = sf.dx = sf.dy = sf.dz = sf.tmax = sf.ex = sf.ey = sf.ez = 0.f;
      if (has) {
        const bool captured = fast_step(bh, fr, sf.ox, sf.oy, sf.oz, sf.dx, sf.dy, sf.dz, sf.tmax);
        sf.ex = e0x; sf.ey = e0y; sf.ez = e0z; e0x = e0y = e0z = 0.f;   /* first chord: leaves from the caller's double origin */
        if (captured) { h.status = 2; h.prim = -1; h.chord = j; h.steps = j + 1; h.t = 0.0; }
        if (COUNT) cnt.steps++;
      }
      const bool live = has && h.status < 0;
      if (W == 2) found = traverse<ANY, COUNT, false, false>(sc, sf, best, stack, RRT_BLOCK, cnt, live, coop);
      else found = traverse_wide<ANY, COUNT, (W == 2 ? 4 : W), false, false>(sc, sf, best, stack, RRT_BLOCK, cnt, live, coop);
      if (live) {
        if (found) {
          h.status = 1; h.prim = best.prim; h.chord = j; h.steps = j + 1; h.t = best.t;
          if (IO::GEO) fill_geo(sc, seg_exact(sf), best, h);
        } else {
          ++j;
          if (j >= bh.nchords) { h.status = 0; h.prim = -1; h.chord = j - 1; h.steps = j; h.t = 0.0; }
        }
      }
    } else if constexpr (SMEM) {
      SegS ss; ss.p = sm;
      if (has) {
        Seg e = seg_exact(ss);
        next_micro_ray(bh, e.o, e.d, e.tmax);
        sm[0] = e.o.x; sm[RRT_SEG_STRIDE] = e.o.y; sm[2 * RRT_SEG_STRIDE] = e.o.z; sm[3 * RRT_SEG_STRIDE] = e.d.x; sm[4 * RRT_SEG_STRIDE] = e.d.y; sm[5 * RRT_SEG_STRIDE] = e.d.z;
        sm[6 * RRT_SEG_STRIDE] = e.tmax;
        if (horizon_hit(bh, e.o, e.d, e.tmax)) { h.status = 2; h.prim = -1; h.chord = j; h.steps = j + 1; h.t = 0.0; }
        if (COUNT) cnt.steps++;
      }
      const bool live = has && h.status < 0;
      bool root_miss = false;
      if (W == 2) found = traverse<ANY, COUNT, false, true, SegS>(sc, ss, best, stack, RRT_BLOCK, cnt, live, coop, &root_miss);
      else found = traverse_wide<ANY, COUNT, (W == 2 ? 4 : W), false, true, SegS>(sc, ss, best, stack, RRT_BLOCK, cnt, live, coop, &root_miss);
      if (live) {
        if (found) {
          h.status = 1; h.prim = best.prim; h.chord = j; h.steps = j + 1; h.t = best.t;
          if (IO::GEO) fill_geo(sc, seg_exact(ss), best, h);
        } else {
          ++j;
          if (j >= bh.nchords) { h.status = 0; h.prim = -1; h.chord = j - 1; h.steps = j; h.t = 0.0; }
        }
        out_streak = root_miss ? out_streak + 1 : 0;
      }
      if constexpr (EW) {
        /* hand a ray that has left the scene box over to warp 0 (see EscQ): not from warp 0 itself, not once the launch's queue is exhausted */
        if (!esc_warp && more && has && h.status < 0 && out_streak >= 2) {
          const int c = atomicAdd(&eq.count, 1);
          if (c >= RRT_EQ_CAP) atomicSub(&eq.count, 1);          /* ring full: the ray stays here */
          else {
            const int q = atomicAdd(&eq.tail, 1) % RRT_EQ_CAP;
#pragma unroll
            for (int cc = 0; cc < 7; ++cc) ((volatile double*)eq.st[cc])[q] = sm[cc * RRT_SEG_STRIDE];
            ((volatile int*)eq.idx)[q] = idx; ((volatile int*)eq.j)[q] = j;
            __threadfence_block();
            ((volatile int*)eq.flag)[q] = 1;
            has = false;
          }
        }
      }
    } else {
      Seg s; s.o = o; s.d = d; s.tmin = tmin; s.tmax = tmax;
      bool parked = false;
      if (has && !(PARK == 2 && skip)) {
        if (!flat) {
          next_micro_ray(bh, o, d, len);
          s.o = o; s.d = d; s.tmin = 0.0; s.tmax = len;
          if (horizon_hit(bh, o, d, len)) { h.status = 2; h.prim = -1; h.chord = j; h.steps = j + 1; h.t = 0.0; }
          else if (PARK == 1 && chord_outside_root(sc, o, len)) {   /* this chord reaches nothing: on to the next one, in k_escape */
            ++j;
            if (j >= bh.nchords) { h.status = 0; h.prim = -1; h.chord = j - 1; h.steps = j; h.t = 0.0; }
            else parked = true;
          }
        }
        if (COUNT) cnt.steps++;
      }
      skip = false;
      if constexpr (PARK == 1) {
        const unsigned pm = __ballot_sync(0xffffffffu, parked);
        if (pm) {
          int base = 0;
          if (wt_lane() == __ffs(pm) - 1) base = atomicAdd(pk.cnt, __popc(pm));
          base = __shfl_sync(0xffffffffu, base, __ffs(pm) - 1);
          if (parked) {
            const size_t q = (size_t)(base + __popc(pm & lt_mask));
            pk.st[q] = o.x; pk.st[pk.cap + q] = o.y; pk.st[2 * pk.cap + q] = o.z; pk.st[3 * pk.cap + q] = d.x; pk.st[4 * pk.cap + q] = d.y; pk.st[5 * pk.cap + q] = d.z;
            pk.st[6 * pk.cap + q] = len; pk.meta[q] = idx; pk.meta[pk.cap + q] = j;
            has = false;
          }
        }
      }
      const bool want = has && h.status < 0 && !chord_is_clear(sc, s.o, s.tmax);
      const bool live = has && h.status < 0;
      /* whole warp, converged */
      if (W == 2) found = traverse<ANY, COUNT, COOP>(sc, s, best, stack, RRT_BLOCK, cnt, want, coop);
      else found = traverse_wide<ANY, COUNT, (W == 2 ? 4 : W), COOP>(sc, s, best, stack, RRT_BLOCK, cnt, want, coop);
      if (live) {
        if (found) {
          h.status = 1; h.prim = best.prim; h.chord = j; h.steps = j + 1; h.t = best.t;
          if (IO::GEO) fill_geo(sc, s, best, h);
        } else {
          ++j;
          if (flat || j >= bh.nchords) { h.status = 0; h.prim = -1; h.chord = flat ? 0 : j - 1; h.steps = flat ? 1 : j; h.t = 0.0; }
        }
      }
    }
    if (has && h.status >= 0) { io.store(idx, h); has = false; }
    __syncwarp();
  }
  if constexpr (EW) {
    if (!esc_warp && wt_lane() == 0) { __threadfence_block(); atomicSub(&eq.producers, 1); }
  }
  if (COUNT) wt_flush_counters(cnt, gcnt);
}

/* the lean loop that carries parked rays: see ParkQ above */
template <bool COUNT, class IO>
__device__ __forceinline__ void escape_persistent(const DevScene& sc, const DevHole& bh, const IO& io, const ParkQ in, const ParkQ out, Counters* gcnt) {
  Counters cnt = {0, 0, 0, 0, 0, 0};
  const int n = in.cnt[0];
  int* cursor = in.cnt + 1;
  const unsigned lt_mask = (1u << wt_lane()) - 1u;
  int pool_next = 0, pool_end = 0;
  bool more = true, has = false;
  int idx = 0, j = 0;
  d3 o = D3(0, 0, 0), d = D3(0, 0, 0);
  double len = 0.0;
  while (true) {
    unsigned need = __ballot_sync(0xffffffffu, !has);
    while (need) {
      if (pool_next >= pool_end) {
        if (!more) break;
        int base = 0;
        if (wt_lane() == 0) base = atomicAdd(cursor, RRT_POOL);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) { more = false; break; }
        pool_next = base; pool_end = min(base + RRT_POOL, n);
      }
      const int avail = pool_end - pool_next;
      const int rank = __popc(need & lt_mask);
      if (!has && rank < avail) {
        const size_t q = (size_t)(pool_next + rank);
        o = D3(in.st[q], in.st[in.cap + q], in.st[2 * in.cap + q]); d = D3(in.st[3 * in.cap + q], in.st[4 * in.cap + q], in.st[5 * in.cap + q]);
        len = in.st[6 * in.cap + q]; idx = in.meta[q]; j = in.meta[in.cap + q];
        has = true;
      }
      pool_next += min(__popc(need), avail);
      need = __ballot_sync(0xffffffffu, !has);
    }
    if (!__any_sync(0xffffffffu, has)) break;
    Hit h; h.status = -1;
    bool reenter = false;
    if (has) {
      next_micro_ray(bh, o, d, len);
      if (COUNT) cnt.steps++;
      if (horizon_hit(bh, o, d, len)) { h.status = 2; h.prim = -1; h.chord = j; h.steps = j + 1; h.t = 0.0; }
      else {
        bool clear = chord_outside_root(sc, o, len);
        if (!clear) {   /* the test traverse() starts with: a chord that fails it visits no node */
          Seg s; s.o = o; s.d = d; s.tmin = 0.0; s.tmax = len;
          BoxCtx bc; float tmaxf, tn;
          seg_boxctx(sc, s, bc, tmaxf);
          clear = !box_hit(sc.root_lo[0], sc.root_lo[1], sc.root_lo[2], sc.root_hi[0], sc.root_hi[1], sc.root_hi[2], bc, tmaxf, tn);
        }
        if (clear) {
          ++j;
          if (j >= bh.nchords) { h.status = 0; h.prim = -1; h.chord = j - 1; h.steps = j; h.t = 0.0; }
        } else reenter = true;
      }
    }
    const unsigned rm = __ballot_sync(0xffffffffu, reenter);
    if (rm) {
      int base = 0;
      if (wt_lane() == __ffs(rm) - 1) base = atomicAdd(out.cnt, __popc(rm));
      base = __shfl_sync(0xffffffffu, base, __ffs(rm) - 1);
      if (reenter) {
        const size_t q = (size_t)(base + __popc(rm & lt_mask));
        out.st[q] = o.x; out.st[out.cap + q] = o.y; out.st[2 * out.cap + q] = o.z; out.st[3 * out.cap + q] = d.x; out.st[4 * out.cap + q] = d.y; out.st[5 * out.cap + q] = d.z;
        out.st[6 * out.cap + q] = len; out.meta[q] = idx; out.meta[out.cap + q] = j;
        has = false;
      }
    }
    if (has && h.status >= 0) { io.store(idx, h); has = false; }
  }
  if (COUNT) wt_flush_counters(cnt, gcnt);
}

#if RRT_AB
/* ---------------------------------------------------------------- STEP = 3: decoupled stepping (hole on, FP64 stepper; RRT_OPT_TRACE_KERNEL = 2, A/B builds).
 * MEASURED SLOWER than the lockstep loop on B200 (round 2, profiles/r02_a_ab_decoupled.jsonl; ms per frame, lockstep -> decoupled -> decoupled + 128-cell grid):
 *   bunny 480x360 64 spp 211.2 -> 217.8 -> 221.8 (node visits 20.2 G -> 20.2 G -> 9.0 G), C2 33.4 -> 35.3 -> 42.3, bunny 1080p 16 spp 97.6 -> 103.2 -> 106.5,
 *   C5 63.3 -> 68.5.  Halving the node visits buys nothing: a warp's traversal lasts as long as its longest lane, and the chords that
 *   still need the BVH are the long ones; the clear tests cost FP64-pipe slots (6 FP64 ops + 4 conversions per chord) the stepper needs.
 * The lockstep loop above pays one BVH traversal per loop trip as long as ANY lane's chord needs one, and the whole warp
 * waits for it.  Most chords need none: a chord that starts outside the scene box and is shorter than its distance to
 * the box (chord_outside_root: the escaped rays of a wide frame, 92 % of a 1080p frame's camera rays), or one that is
 * shorter than the clearance of the grid cell it starts in (chord_is_clear: open space inside the box), provably meets
 * no primitive, so BVHAccel::intersect_micro would return false for it (reference src/bvh.cpp:107-110 only ever looks at
 * primitives the chord reaches).  Here a lane whose chord is provably clear simply steps on; a lane whose chord does need
 * the BVH parks it (`pend`), and the warp votes each trip: STEP while at least RRT_S_THRESH lanes can still step (or
 * nobody is parked), otherwise TRAVERSE the parked chords — with most lanes taking part instead of a few.  Arithmetic of
 * every chord, horizon test and primitive test is untouched and chords are visited in order, so results are bit-identical
 * to the lockstep loop (test_bvh_intersect_matches_reference / test_far_origin_chords_match_oracle run it in A/B builds). */
#ifndef RRT_S_THRESH
#define RRT_S_THRESH 8
#endif
template <bool ANY, bool COUNT, int W, class IO>
__device__ __forceinline__ void trace_decoupled(const DevScene& sc, const DevHole& bh, const IO& io, const int n, int* cursor, Counters* gcnt) {
  __shared__ int stack_mem[RRT_STACK * RRT_BLOCK];
  int* stack = stack_mem + threadIdx.x;
  Counters cnt = {0, 0, 0, 0, 0, 0};
  const unsigned lt_mask = (1u << wt_lane()) - 1u;
  int pool_next = 0, pool_end = 0;   /* warp-uniform */
  bool more = true;                  /* warp-uniform */
  bool has = false, pend = false;    /* pend: the current chord (o, d, len) waits for its BVH traversal */
  int idx = 0, j = 0;
  d3 o = D3(0, 0, 0), d = D3(0, 0, 0);
  double len = 0.0;
  while (true) {
    unsigned need = __ballot_sync(0xffffffffu, !has);
    while (need) {
      if (pool_next >= pool_end) {
        if (!more) break;
        int base = 0;
        if (wt_lane() == 0) base = atomicAdd(cursor, RRT_POOL);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) { more = false; break; }
        pool_next = base; pool_end = min(base + RRT_POOL, n);
      }
      const int avail = pool_end - pool_next;
      const int rank = __popc(need & lt_mask);
      if (!has && rank < avail) {
        double tmin_unused, tmax_unused;   /* the bender discards the caller's range (reference src/bvh.cpp:104) */
        idx = pool_next + rank;
        io.load(idx, o, d, tmin_unused, tmax_unused);
        has = true; pend = false; j = 0; len = 0.0;
        if (COUNT) cnt.rays++;
      }
      pool_next += min(__popc(need), avail);
      need = __ballot_sync(0xffffffffu, !has);
    }
    const unsigned m_has = __ballot_sync(0xffffffffu, has);
    if (!m_has) break;
    const unsigned m_step = __ballot_sync(0xffffffffu, has && !pend);
    Hit h; h.status = -1;
    if (__popc(m_step) >= RRT_S_THRESH || m_step == m_has) {
      /* ---- STEP: one chord for every lane that is not parked */
      if (has && !pend) {
        next_micro_ray(bh, o, d, len);
        if (COUNT) cnt.steps++;
        if (horizon_hit(bh, o, d, len)) { h.status = 2; h.prim = -1; h.chord = j; h.steps = j + 1; h.t = 0.0; }
        else if (chord_is_clear(sc, o, len) || chord_outside_root(sc, o, len)) {
          ++j;
          if (j >= bh.nchords) { h.status = 0; h.prim = -1; h.chord = j - 1; h.steps = j; h.t = 0.0; }
        } else pend = true;
      }
    } else {
      /* ---- TRAVERSE: the parked chords, whole warp converged */
      Seg s; s.o = o; s.d = d; s.tmin = 0.0; s.tmax = len;
      Best best;
      bool found;
      if (W == 2) found = traverse<ANY, COUNT, false>(sc, s, best, stack, RRT_BLOCK, cnt, pend, nullptr);
      else found = traverse_wide<ANY, COUNT, (W == 2 ? 4 : W), false>(sc, s, best, stack, RRT_BLOCK, cnt, pend, nullptr);
      if (pend) {
        pend = false;
        if (found) {
          h.status = 1; h.prim = best.prim; h.chord = j; h.steps = j + 1; h.t = best.t;
          if (IO::GEO) fill_geo(sc, s, best, h);
        } else {
          ++j;
          if (j >= bh.nchords) { h.status = 0; h.prim = -1; h.chord = j - 1; h.steps = j; h.t = 0.0; }
        }
      }
    }
    if (has && h.status >= 0) { io.store(idx, h); has = false; }
    __syncwarp();
  }
  if (COUNT) wt_flush_counters(cnt, gcnt);
}

/* ---------------------------------------------------------------- variant 1: vote-scheduled state machine (measured
 * SLOWER than variant 0 on C2, 62.9 vs 48.2 ms per frame: kept for A/B).
 * The lockstep loop above still loses lanes INSIDE a chord: traversals have different lengths and lanes that reach a
 * leaf run the FP64 primitive test while their neighbours wait (ncu on variant 0: 13-18 active threads per
 * instruction, DMUL/DADD at 13-14).  Here every lane is in one of three work states — STEP (advance the geodesic by
 * one chord and set up its traversal), NODE (visit one interior node), LEAF (test one leaf record) — and each trip
 * of the loop the warp votes and executes only the block most lanes are waiting for.  A block therefore never runs
 * for fewer than a third of the lanes that have work, and the expensive FP64 blocks run nearly full. */
template <bool ANY, bool COUNT, class IO>
__device__ __forceinline__ void trace_sched(const DevScene& sc, const DevHole& bh, const IO& io, const int n, int* cursor, Counters* gcnt) {
  __shared__ int stack_mem[RRT_STACK * RRT_BLOCK];
  int* stack = stack_mem + threadIdx.x;
  enum { ST_IDLE = 0, ST_FREE = 1, ST_STEP = 2, ST_NODE = 3, ST_LEAF = 4 };
  Counters cnt = {0, 0, 0, 0, 0, 0};
  const bool flat = bh.mode == RRT_MODE_FLAT;
  const unsigned lt_mask = (1u << wt_lane()) - 1u;
  int pool_next = 0, pool_end = 0;
  bool more = true;
  int state = ST_FREE;
  int idx = 0, j = 0, cur = 0, sp = 0;
  uint32_t lf_next = 0, lf_end = 0;
  bool found = false, shifted = false;
  float tmaxf = 0.f;
  Seg s; s.o = s.d = D3(0, 0, 0); s.tmin = s.tmax = 0.0;
  Best best; best.t = 0.0; best.b1 = best.b2 = 0.0; best.prim = -1; best.rank = -1; best.kind = 0;
  BoxCtx bc; bc.ix = bc.iy = bc.iz = bc.nx_ = bc.ny_ = bc.nz_ = bc.fx_ = bc.fy_ = bc.fz_ = bc.tmin = 0.f; bc.nx = bc.ny = bc.nz = false;

  auto finish = [&](int status) {
    Hit h; h.status = status; h.prim = status == 1 ? best.prim : -1; h.t = status == 1 ? best.t : 0.0;
    if (status == 0) { h.chord = flat ? 0 : j - 1; h.steps = flat ? 1 : j; }
    else { h.chord = j; h.steps = j + 1; }
    if (IO::GEO && status == 1) fill_geo(sc, s, best, h);
    io.store(idx, h);
    state = ST_FREE;
  };
  auto chord_done = [&]() {   /* traversal of the current chord is exhausted */
    if (found) finish(1);
    else { ++j; if (flat || j >= bh.nchords) finish(0); else state = ST_STEP; }
  };
  auto enter = [&](int node) {  /* next thing to visit */
    cur = node;
    if (node < 0) { uint32_t bits = (uint32_t)~node; lf_next = bits >> 3; lf_end = lf_next + (bits & 7u); state = lf_next < lf_end ? ST_LEAF : ST_NODE; }
    else state = ST_NODE;
  };
  auto pop_or_done = [&]() {
    if (sp == 0) chord_done();
    else enter(stack[(--sp) * RRT_BLOCK]);
  };

  while (true) {
    /* ---- refill free lanes from the warp pool */
    unsigned need = __ballot_sync(0xffffffffu, state == ST_FREE);
    while (need) {
      if (pool_next >= pool_end) {
        if (!more) break;
        int base = 0;
        if (wt_lane() == 0) base = atomicAdd(cursor, RRT_POOL);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) { more = false; break; }
        pool_next = base; pool_end = min(base + RRT_POOL, n);
      }
      const int avail = pool_end - pool_next;
      const int rank = __popc(need & lt_mask);
      if (state == ST_FREE && rank < avail) {
        idx = pool_next + rank;
        io.load(idx, s.o, s.d, s.tmin, s.tmax);
        if (!flat) { s.tmin = 0.0; s.tmax = 0.0; }
        j = 0; state = ST_STEP;
        if (COUNT) cnt.rays++;
      }
      pool_next += min(__popc(need), avail);
      need = __ballot_sync(0xffffffffu, state == ST_FREE);
    }
    if (state == ST_FREE) state = ST_IDLE;   /* queue exhausted */
    /* ---- vote */
    const int nS = __popc(__ballot_sync(0xffffffffu, state == ST_STEP));
    const int nN = __popc(__ballot_sync(0xffffffffu, state == ST_NODE));
    const int nL = __popc(__ballot_sync(0xffffffffu, state == ST_LEAF));
    if ((nS | nN | nL) == 0) break;
    if (nN >= nL && nN >= nS) {
      if (state == ST_NODE) {
        if (cur < 0) pop_or_done();   /* empty leaf */
        else {
          const float4* nd = sc.nodes + 4 * (size_t)cur;
          float4 q0 = __ldg(nd), q1 = __ldg(nd + 1), q2 = __ldg(nd + 2), q3 = __ldg(nd + 3);
          if (COUNT) cnt.nodes++;
          float tl, tr;
          bool hl = box_hit(q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, bc, tmaxf, tl);
          bool hr = box_hit(q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, bc, tmaxf, tr);
          int cl = __float_as_int(q3.x), cr = __float_as_int(q3.y);
          if (hl && hr) {
            if (tr < tl) { int t = cl; cl = cr; cr = t; }
            stack[(sp++) * RRT_BLOCK] = cr;
            enter(cl);
          } else if (hl) enter(cl);
          else if (hr) enter(cr);
          else pop_or_done();
        }
      }
    } else if (nL >= nS) {
      if (state == ST_LEAF) {
        bool acc = leaf_record_test<COUNT>(sc.leaf + 5 * (size_t)lf_next, s, best, cnt);
        ++lf_next;
        if (acc) {
          found = true;
          if (ANY) finish(1);
          else if (!shifted) tmaxf = __double2float_ru(best.t);
        }
        if (state == ST_LEAF && lf_next >= lf_end) pop_or_done();
      }
    } else {
      if (state == ST_STEP) {
        bool captured = false;
        if (!flat) {
          next_micro_ray(bh, s.o, s.d, s.tmax);
          captured = horizon_hit(bh, s.o, s.d, s.tmax);
        }
        if (COUNT) cnt.steps++;
        if (captured) finish(2);
        else {
          best.t = s.tmax; best.rank = -1; best.prim = -1; best.kind = 0; best.b1 = best.b2 = 0.0;
          found = false; sp = 0;
          tmaxf = __double2float_ru(best.t);
          float tr;
          if (chord_is_clear(sc, s.o, s.tmax)) chord_done();
          else {
            const double T = boxctx_make(sc, s.o, s.d, s.tmin, bc);
            shifted = T != 0.0;
            tmaxf = boxctx_tmax(best.t, T);
            if (box_hit(sc.root_lo[0], sc.root_lo[1], sc.root_lo[2], sc.root_hi[0], sc.root_hi[1], sc.root_hi[2], bc, tmaxf, tr)) enter(sc.root);
            else chord_done();
          }
        }
      }
    }
    __syncwarp();
  }
  if (COUNT) wt_flush_counters(cnt, gcnt);
}

#endif  /* RRT_AB */

/* extension rays: closest hit */
#ifdef RRT_MINB_CLOSEST_N      /* A/B: resident blocks for the closest-hit kernel alone */
#define RRT_MINB_CLOSEST(V, STEP) RRT_MINB_CLOSEST_N
#else   /* wide closest-hit kernel: 5 blocks at 96 registers (fewer spills around Best / fill_geo): C2 37.08 -> 36.51 ms, C1 4.47 -> 4.42;
           the binary one stays at 6 (5: bunny 1080p 103.8 -> 104.8, C5 68.6 -> 70.4) */
#ifndef RRT_MINB_CLOSEST_BIN
#define RRT_MINB_CLOSEST_BIN RRT_MINB   /* A/B: resident blocks of the binary-node closest-hit kernel alone */
#endif
#define RRT_MINB_CLOSEST(V, STEP) (((V) >= 4 && (STEP) != 1) ? 5 : ((V) < 4 && (STEP) != 1) ? RRT_MINB_CLOSEST_BIN : RRT_MINBLOCKS(V, STEP))
#endif
template <bool COUNT, int V, int STEP = 0, int PARK = 0>
__global__ void __launch_bounds__(RRT_BLOCK, RRT_MINB_CLOSEST(V, STEP)) k_trace_closest(const __grid_constant__ DevScene sc, const __grid_constant__ DevHole bh,
                                                                const double* __restrict__ ray, size_t cap, const int* __restrict__ count_ptr,
                                                                int* cursor, int* __restrict__ h_prim, double* __restrict__ h_geo, Counters* gcnt, const ParkQ pk) {
  ClosestIO io; io.ray = ray; io.cap = cap; io.h_prim = h_prim; io.h_geo = h_geo;
#if RRT_AB
  if constexpr (V == 1) trace_sched<false, COUNT>(sc, bh, io, *count_ptr, cursor, gcnt);
  else if constexpr (STEP == 3) trace_decoupled<false, COUNT, (V == 0 ? 2 : V)>(sc, bh, io, *count_ptr, cursor, gcnt);
  else
#endif
  trace_persistent<false, COUNT, (V == 0 ? 2 : V), STEP, ClosestIO, PARK>(sc, bh, io, PARK == 2 ? pk.cnt[0] : *count_ptr, PARK == 2 ? pk.cnt + 1 : cursor, gcnt, pk);
}

/* shadow rays: BVHAccel::intersect(ray) (any hit), or closest hit when the hemisphere estimator needs the emitter */
template <bool ANY, bool COUNT, int V, int STEP = 0, int PARK = 0>
__global__ void __launch_bounds__(RRT_BLOCK, RRT_MINBLOCKS(V, STEP)) k_trace_shadow(const __grid_constant__ DevScene sc, const __grid_constant__ DevHole bh,
                                                               const double* __restrict__ ray, size_t cap, const int* __restrict__ count_ptr,
                                                               int* cursor, const uint32_t* __restrict__ cidx, int* __restrict__ res,
                                                               const uint32_t* __restrict__ perm, Counters* gcnt, const ParkQ pk) {
  ShadowIO io; io.ray = ray; io.cap = cap; io.cidx = cidx; io.res = res; io.perm = perm;
#if RRT_AB
  if constexpr (V == 1) trace_sched<ANY, COUNT>(sc, bh, io, *count_ptr, cursor, gcnt);
  else if constexpr (STEP == 3) trace_decoupled<ANY, COUNT, (V == 0 ? 2 : V)>(sc, bh, io, *count_ptr, cursor, gcnt);
  else
#endif
  trace_persistent<ANY, COUNT, (V == 0 ? 2 : V), STEP, ShadowIO, PARK>(sc, bh, io, PARK == 2 ? pk.cnt[0] : *count_ptr, PARK == 2 ? pk.cnt + 1 : cursor, gcnt, pk);
}

/* parity hook: the same loop on caller-supplied rays */
template <bool ANY, int V, int STEP = 0, int PARK = 0>
__global__ void __launch_bounds__(RRT_BLOCK, RRT_MINBLOCKS(V, STEP)) k_hook_trace(const __grid_constant__ DevScene sc, const __grid_constant__ DevHole bh, int n,
                                                            const double* __restrict__ o, const double* __restrict__ d, const double* __restrict__ tmin,
                                                            const double* __restrict__ tmax, rrt_hit* __restrict__ out, int* cursor, Counters* gcnt, const ParkQ pk) {
  HookIO<ANY> io; io.o = o; io.d = d; io.tmin = tmin; io.tmax = tmax; io.out = out;
#if RRT_AB
  if constexpr (V == 1) trace_sched<ANY, true>(sc, bh, io, n, cursor, gcnt);
  else if constexpr (STEP == 3) trace_decoupled<ANY, true, (V == 0 ? 2 : V)>(sc, bh, io, n, cursor, gcnt);
  else
#endif
  trace_persistent<ANY, true, (V == 0 ? 2 : V), STEP, HookIO<ANY>, PARK>(sc, bh, io, PARK == 2 ? pk.cnt[0] : n, PARK == 2 ? pk.cnt + 1 : cursor, gcnt, pk);
}

/* k_escape: carries the rays the trace kernel parked (ParkQ above).  No BVH, no stack: 58 registers, 8 blocks of 128 threads per SM */
#ifndef RRT_MINB_ESCAPE
#define RRT_MINB_ESCAPE 8
#endif
template <bool COUNT>
__global__ void __launch_bounds__(RRT_BLOCK, RRT_MINB_ESCAPE) k_escape_closest(const __grid_constant__ DevScene sc, const __grid_constant__ DevHole bh, size_t cap,
                                                                               int* __restrict__ h_prim, double* __restrict__ h_geo, Counters* gcnt, const ParkQ in, const ParkQ out) {
  ClosestIO io; io.ray = nullptr; io.cap = cap; io.h_prim = h_prim; io.h_geo = h_geo;
  escape_persistent<COUNT>(sc, bh, io, in, out, gcnt);
}
template <bool COUNT>
__global__ void __launch_bounds__(RRT_BLOCK, RRT_MINB_ESCAPE) k_escape_shadow(const __grid_constant__ DevScene sc, const __grid_constant__ DevHole bh, size_t cap,
                                                                              const uint32_t* __restrict__ cidx, int* __restrict__ res, Counters* gcnt, const ParkQ in, const ParkQ out) {
  ShadowIO io; io.ray = nullptr; io.cap = cap; io.cidx = cidx; io.res = res; io.perm = nullptr;
  escape_persistent<COUNT>(sc, bh, io, in, out, gcnt);
}
template <bool ANY>
__global__ void __launch_bounds__(RRT_BLOCK, RRT_MINB_ESCAPE) k_escape_hook(const __grid_constant__ DevScene sc, const __grid_constant__ DevHole bh, rrt_hit* __restrict__ hits,
                                                                            Counters* gcnt, const ParkQ in, const ParkQ out) {
  HookIO<ANY> io; io.o = io.d = io.tmin = io.tmax = nullptr; io.out = hits;
  escape_persistent<true>(sc, bh, io, in, out, gcnt);
}
