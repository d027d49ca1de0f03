// rrt_kernels.cuh — the wavefront kernels: generate, trace (geodesic-advance + intersect), shade / BSDF-sample,
// accumulate, resolve, plus the stable compaction used for the active-pixel list.
//
// Queues are SoA arrays of capacity `cap` (the stride between components).  A path is identified by its
// SLOT = pixel_in_chunk * samples_this_round + sample_in_round, which never changes; queue POSITIONS change
// every bounce because the shade kernel compacts survivors with warp ballot + one atomic per warp.
#pragma once
#include "rrt_trace.cuh"
#include "rrt_shade.cuh"

#define RRT_BLOCK 128

struct RenderK {      // per-render constants (reference PathTracer members, src/pathtracer.cpp:32-85)
  int W, H;
  int ns_aa, D, ns_area_light, batch;
  float tol;
  int hemi, thin_lens;
  unsigned long long seed;
  int S;              // light samples taken per shaded vertex (slots in the contribution table)
  int total;          // divisor of the direct estimate (reference src/part1_code.cpp:17,56)
  int cap;            // queue capacity = SoA stride
  int direct_draws;   // uniforms one direct-lighting estimate consumes (fixed per scene: 2 per sample of an area / hemisphere / environment light)
  int split;          // 1: direct lighting in k_shade_direct (one thread per light sample), k_shade does emission + bounce only
  int defer_emit;     // 1: k_accumulate runs for every vertex index (D >= 1 and S > 0), so the emission term of a vertex reached through a delta BSDF is
                      //    handed to it (Queues::emit) instead of being added by k_shade: k_shade(it + 1) then touches nothing k_accumulate(it) touches
};

struct Queues {
  double* ray[2];       // 8 x cap: o.xyz d.xyz tmin tmax          (extension rays, ping-pong)
  uint32_t* slot[2];    // path slot
  uint32_t* rngn[2];    // draws consumed so far
  uint32_t* flags[2];   // bit0: previous vertex was a delta BSDF
  float* beta[2];       // 3 x cap: throughput
  int* h_prim;          // closest-hit result per queue position: prim >= 0, -1 miss, -2 captured
  double* h_geo;        // 9 x cap: hit_p, w_out, n
  /* shadow side, double-buffered by the parity of the vertex index `it`: the shadow rays of vertex it are traced (and folded by k_accumulate(it)) on
   * their own stream while k_shade / k_shade_direct already fill the buffers of vertex it + 1 (rrt_api.cu, RRT_OPT_OVERLAP) */
  float* dbeta[2];      // 3 x cap: throughput at the vertex whose direct light is pending
  uint8_t* vflag[2];    // bit0: vertex took light samples; bit1: an emission term waits in `emit` (previous vertex was a delta BSDF)
  float* emit[2];       // 3 x cap: Le * throughput of such a vertex, added by k_accumulate BEFORE the vertex's direct light (the reference's order)
  uint32_t* aslot[2];   // path slot of the vertex, for k_accumulate
  double* sray[2];      // 7 x capS: shadow rays o d tmax
  uint32_t* s_cidx[2];  // contribution-table index of each shadow ray
  int* s_res[2];        // capS: -2 no ray, -1 unoccluded, >= 0 prim hit
  float4* s_contrib[2]; // capS: contribution (rgb), w = cos term (hemisphere mode)
  uint16_t* s_bin[2];   // capS: origin cell of each shadow ray       } RRT_OPT_SORT_SHADOW, k_bin_* below
  uint32_t* s_perm[2];  // capS: launch order -> shadow-queue position }
  uint32_t* s_hist[2];  // 2 x RRT_BINS: rays per cell (zero between uses), first launch position of each cell
  float* Lsample;       // 3 x cap: radiance of each path slot
  int* ctl;             // [4][D+2]: queue counts, shadow counts, trace cursors, shadow cursors
  int ctl_stride;
};

struct PixelAcc {      // per-pixel state across adaptive rounds (reference src/part1_code.cpp:125-163)
  float* sum;          // 3 x npix
  double* s1; double* s2;
  int* n;              // samples taken
  uint8_t* done;       // adaptive stop reached
};

__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

/* warp-aggregated append from possibly divergent code: one atomic per converged group */
__device__ __forceinline__ int append_pos(int* counter) {
  unsigned m = __activemask();
  int leader = __ffs(m) - 1;
  int rank = __popc(m & ((1u << lane_id()) - 1u));
  int base = 0;
  if (lane_id() == leader) base = atomicAdd(counter, __popc(m));
  base = __shfl_sync(m, base, leader);
  return base + rank;
}

__device__ __forceinline__ void flush_counters(const Counters& c, Counters* g) {
  unsigned long long v[5] = {c.rays, c.steps, c.nodes, c.tris, c.sph};
  unsigned long long* dst = (unsigned long long*)g;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    unsigned long long x = v[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if (lane_id() == 0 && x) atomicAdd(dst + k, x);
  }
}

/* ---------------------------------------------------------------- generate: Camera::generate_ray per path slot
 * (reference src/part1_code.cpp:125-145,182-187) */
__global__ void __launch_bounds__(256) k_generate(const __grid_constant__ DevCamera cam, const __grid_constant__ RenderK rk,
                                                  const __grid_constant__ Queues q, const uint32_t* __restrict__ act, int npix, int base_sample, int nb) {
  int npaths = npix * nb;
  for (int slot = blockIdx.x * blockDim.x + threadIdx.x; slot < npaths; slot += gridDim.x * blockDim.x) {
    uint32_t pixel = act[slot / nb];
    uint32_t sample = (uint32_t)(base_sample + slot % nb);
    int x = (int)(pixel % (uint32_t)rk.W), y = (int)(pixel / (uint32_t)rk.W);
    Rng g = rng_make(rk.seed, pixel, sample, 0);
    double sx = x, sy = y;
    if (rk.ns_aa == 1) { sx += 0.5; sy += 0.5; }
    else { double jx, jy; rng_grid2(g, jx, jy); sx += jx; sy += jy; }
    d3 o, d;
    if (rk.thin_lens) {
      double lx, ly; rng_grid2(g, lx, ly);
      generate_ray_thin_lens(cam, sx / rk.W, sy / rk.H, lx, ly * 2 * RRT_PI, o, d);
    } else generate_ray(cam, sx / rk.W, sy / rk.H, o, d);
    double* r = q.ray[0];
    size_t c = (size_t)rk.cap;
    r[slot] = o.x; r[c + slot] = o.y; r[2 * c + slot] = o.z; r[3 * c + slot] = d.x; r[4 * c + slot] = d.y; r[5 * c + slot] = d.z;
    r[6 * c + slot] = cam.nClip; r[7 * c + slot] = cam.fClip;
    q.slot[0][slot] = (uint32_t)slot; q.rngn[0][slot] = g.n; q.flags[0][slot] = 0;
    q.beta[0][slot] = 1.f; q.beta[0][c + slot] = 1.f; q.beta[0][2 * c + slot] = 1.f;
    q.Lsample[slot] = 0.f; q.Lsample[c + slot] = 0.f; q.Lsample[2 * c + slot] = 0.f;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) q.ctl[0] = npaths;
}

#include "rrt_wavetrace.cuh"

/* ---------------------------------------------------------------- shade: est_radiance_global_illumination /
 * at_least_one_bounce_radiance / estimate_direct_lighting_* unrolled over the wavefront
 * (reference src/part1_code.cpp:15-123).  `it` = vertex index along the path (0 = camera hit); the reference's
 * recursion depth at this vertex is D - it. */
#ifndef RRT_SHADE_MINB
#define RRT_SHADE_MINB 4   /* measured on C1: 3 / 4 / 5 / 6 blocks -> 4.87 / 4.43 / 4.69 / 4.75 ms per frame */
#endif
__global__ void __launch_bounds__(RRT_BLOCK, RRT_SHADE_MINB) k_shade(const __grid_constant__ DevScene sc, const __grid_constant__ DevHole bh,
                                                     const __grid_constant__ RenderK rk, const __grid_constant__ Queues q,
                                                     const uint32_t* __restrict__ act, int base_sample, int nb, int it) {
  const int cur = it & 1, nxt = cur ^ 1;
  const int sp = it & 1;            /* parity of the shadow-side buffers */
  const size_t cap = (size_t)rk.cap;
  const int count = q.ctl[it];
  int* next_count = q.ctl + (it + 1);
  int* sh_count = q.ctl + q.ctl_stride + it;
  uint8_t* const vflag = q.vflag[sp]; float* const dbeta = q.dbeta[sp];
  double* const sray = q.sray[sp]; uint32_t* const s_cidx = q.s_cidx[sp]; int* const s_res = q.s_res[sp]; float4* const s_contrib = q.s_contrib[sp];
  const int nwarps = (gridDim.x * blockDim.x) >> 5, warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  for (int base = warp * 32; base < count; base += nwarps * 32) {
    const int i = base + lane_id();
    bool ext = false;
    uint8_t vf_out = 0;
    d3 eo = D3(0, 0, 0), ed = D3(0, 0, 0);
    f3 ebeta = F3(0, 0, 0);
    uint32_t slot = 0, en = 0, eflags = 0;
    if (i < count) {
      const int prim = q.h_prim[i];
      if (prim < 0 && it == 0 && sc.env) {
        /* camera ray that hit nothing (or fell into the hole): the environment seen along its own, unbent direction
         * (reference src/part1_code.cpp:106-107) */
        const double* r = q.ray[cur];
        const f3 Le = env_sample_dir(sc, D3(r[3 * cap + i], r[4 * cap + i], r[5 * cap + i]));
        const uint32_t sl = q.slot[cur][i];
        q.Lsample[sl] = Le.r; q.Lsample[cap + sl] = Le.g; q.Lsample[2 * cap + sl] = Le.b;
      }
      if (prim >= 0) {
        slot = q.slot[cur][i];
        const uint32_t flags = q.flags[cur][i];
        const f3 beta = F3(q.beta[cur][i], q.beta[cur][cap + i], q.beta[cur][2 * cap + i]);
        const rrt_material m = sc.mats[sc.prim_mat[prim]];
        const f3 Le = mat_emission(m);
        if (it == 0) { q.Lsample[slot] = Le.r; q.Lsample[cap + slot] = Le.g; q.Lsample[2 * cap + slot] = Le.b; }
        else if (flags & 1u) {
          f3 a = s_mul(Le, beta);
          if (rk.defer_emit) { vf_out |= 2; q.emit[sp][i] = a.r; q.emit[sp][cap + i] = a.g; q.emit[sp][2 * cap + i] = a.b; }
          else { q.Lsample[slot] += a.r; q.Lsample[cap + slot] += a.g; q.Lsample[2 * cap + slot] += a.b; }
        }
        q.aslot[sp][i] = slot;
        if (rk.D >= 1) {
          const uint32_t pixel = act[slot / (uint32_t)nb];
          Rng g = rng_make(rk.seed, pixel, (uint32_t)(base_sample + slot % (uint32_t)nb), q.rngn[cur][i]);
          const d3 hit_p = D3(q.h_geo[i], q.h_geo[cap + i], q.h_geo[2 * cap + i]);
          const d3 wo_world = D3(q.h_geo[3 * cap + i], q.h_geo[4 * cap + i], q.h_geo[5 * cap + i]);
          const d3 nrm = D3(q.h_geo[6 * cap + i], q.h_geo[7 * cap + i], q.h_geo[8 * cap + i]);
          d3 o2w[3]; make_coord_space(o2w, nrm);
          const d3 w_out = mat_tmul(o2w, wo_world);
          const bool delta = mat_is_delta(m);
          if (rk.D == 1 || !delta) {
            /* one_bounce_radiance: light samples -> shadow queue, contributions -> table row i */
            vf_out |= 1;
            dbeta[i] = beta.r; dbeta[cap + i] = beta.g; dbeta[2 * cap + i] = beta.b;
            int k = 0;
            if (rk.split) {
              g.n += (uint32_t)rk.direct_draws;   /* k_shade_direct takes these draws, one thread per sample */
            } else if (rk.hemi) {
              /* estimate_direct_lighting_hemisphere (reference src/part1_code.cpp:15-31) */
              for (; k < rk.S; ++k) {
                d3 w_in = hemisphere_sample(g);
                d3 wi_world = mat_mul(o2w, w_in);
                f3 f = bsdf_f(m, w_out, w_in);
                const size_t ci = (size_t)i * rk.S + k;
                s_contrib[ci] = make_float4(f.r, f.g, f.b, (float)w_in.z);
                s_res[ci] = -1;
                int pos = append_pos(sh_count);
                const size_t cs = cap * (size_t)rk.S;
                d3 so = vadd(hit_p, smul(RRT_EPS_D, wi_world));
                sray[pos] = so.x; sray[cs + pos] = so.y; sray[2 * cs + pos] = so.z;
                sray[3 * cs + pos] = wi_world.x; sray[4 * cs + pos] = wi_world.y; sray[5 * cs + pos] = wi_world.z;
                sray[6 * cs + pos] = INFINITY;
                s_cidx[pos] = (uint32_t)ci;
              }
            } else {
              /* estimate_direct_lighting_importance (reference src/part1_code.cpp:33-57) */
              const int n_all = sc.n_lights + (sc.env ? 1 : 0);   /* the environment light is the last light (reference src/pathtracer.cpp:107-109) */
              for (int li = 0; li < n_all; ++li) {
                const bool is_env = li >= sc.n_lights;
                const rrt_light& l = sc.lights[is_env ? 0 : li];
                const int ns = (!is_env && l.is_delta) ? 1 : rk.ns_area_light;
                for (int s = 0; s < ns; ++s, ++k) {
                  d3 wi_world; float dist, pdf;
                  f3 Li = is_env ? env_sample_L(sc, wi_world, dist, pdf, g) : light_sample_L(l, hit_p, wi_world, dist, pdf, g);
                  const size_t ci = (size_t)i * rk.S + k;
                  s_res[ci] = -2;
                  if (!is_env && l.type == RRT_LIGHT_NULL) continue;
                  d3 w_in = mat_tmul(o2w, wi_world);
                  if (w_in.z < 0) continue;
                  f3 c = s_divf(s_mulf(s_mul(Li, bsdf_f(m, w_out, w_in)), (float)w_in.z), pdf);
                  s_contrib[ci] = make_float4(c.r, c.g, c.b, 0.f);
                  s_res[ci] = -1;
                  int pos = append_pos(sh_count);
                  const size_t cs = cap * (size_t)rk.S;
                  d3 so = vadd(hit_p, smul(RRT_EPS_D, wi_world));
                  sray[pos] = so.x; sray[cs + pos] = so.y; sray[2 * cs + pos] = so.z;
                  sray[3 * cs + pos] = wi_world.x; sray[4 * cs + pos] = wi_world.y; sray[5 * cs + pos] = wi_world.z;
                  sray[6 * cs + pos] = (double)dist;
                  s_cidx[pos] = (uint32_t)ci;
                }
              }
            }
          }
          if (rk.D >= 2) {
            /* at_least_one_bounce_radiance (reference src/part1_code.cpp:82-97): Russian roulette, BSDF sample */
            const int depth = rk.D - it;
            const double prob = 0.7;
            if (depth == rk.D || (depth > 1 && rng_uniform(g) < prob)) {
              d3 w_in; float pdf;
              f3 f = bsdf_sample_f(m, w_out, w_in, pdf, g);
              if (pdf != 0.0f) {
                ed = mat_mul(o2w, w_in);
                eo = vadd(hit_p, smul(RRT_EPS_D, ed));
                ebeta = s_divf(s_divf(s_mulf(s_mul(beta, f), (float)fabs(w_in.z)), pdf), (float)prob);
                en = g.n; eflags = delta ? 1u : 0u;
                ext = true;
              }
            }
          }
        }
      }
    }
    if (i < count) vflag[i] = vf_out;
    __syncwarp();
    /* compaction of survivors: warp ballot + prefix, one atomic per warp */
    unsigned mask = __ballot_sync(0xffffffffu, ext);
    if (mask) {
      int wbase = 0;
      if (lane_id() == 0) wbase = atomicAdd(next_count, __popc(mask));
      wbase = __shfl_sync(0xffffffffu, wbase, 0);
      if (ext) {
        int pos = wbase + __popc(mask & ((1u << lane_id()) - 1u));
        double* r = q.ray[nxt];
        r[pos] = eo.x; r[cap + pos] = eo.y; r[2 * cap + pos] = eo.z; r[3 * cap + pos] = ed.x; r[4 * cap + pos] = ed.y; r[5 * cap + pos] = ed.z;
        r[6 * cap + pos] = 0.0; r[7 * cap + pos] = INFINITY;
        q.slot[nxt][pos] = slot; q.rngn[nxt][pos] = en; q.flags[nxt][pos] = eflags;
        q.beta[nxt][pos] = ebeta.r; q.beta[nxt][cap + pos] = ebeta.g; q.beta[nxt][2 * cap + pos] = ebeta.b;
      }
    }
  }
}

/* ---------------------------------------------------------------- direct lighting, one thread per (vertex, light sample).
 * The same estimate k_shade makes in its light loop (estimate_direct_lighting_importance / _hemisphere, reference
 * src/part1_code.cpp:15-57), spread over S times as many threads: a sample's uniforms sit at a fixed offset of the
 * vertex's draw counter (every light class takes a fixed number of draws per sample and the RNG is counter based),
 * so sample k needs nothing from samples 0..k-1.  Results are addressed by (vertex, sample) as before; the order in
 * which shadow rays enter their queue changes, the image does not. */
#ifndef RRT_DIRECT_MINB
#define RRT_DIRECT_MINB 4   /* measured (C1 / C3 at 32 spp, ms per frame): 2 -> 5.22 / 27.7, 3 -> 4.82 / 27.0, 4 -> 4.74 / 26.6, 6 -> 4.92 / 26.7, 8 -> 5.89 / 27.7 */
#endif
__global__ void __launch_bounds__(256, RRT_DIRECT_MINB) k_shade_direct(const __grid_constant__ DevScene sc, const __grid_constant__ RenderK rk,
                                                      const __grid_constant__ Queues q, const uint32_t* __restrict__ act, int base_sample, int nb, int it) {
  const int cur = it & 1, sp = it & 1;
  const size_t cap = (size_t)rk.cap;
  double* const sray = q.sray[sp]; uint32_t* const s_cidx = q.s_cidx[sp]; int* const s_res = q.s_res[sp]; float4* const s_contrib = q.s_contrib[sp];
  const int count = q.ctl[it];
  int* sh_count = q.ctl + q.ctl_stride + it;
  const int S = rk.S;
  const long long total = (long long)count * S;
  const long long span = (total + 31) & ~31LL;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < span; t += (long long)gridDim.x * blockDim.x) {
    bool emit = false;
    d3 so = D3(0, 0, 0), wi_world = D3(0, 0, 0);
    double dist_d = 0.0;
    size_t ci = 0;
    if (t < total) {
      const int i = (int)(t / S), k = (int)(t - (long long)i * S);
      const int prim = q.h_prim[i];
      if (prim >= 0) {
        const rrt_material m = sc.mats[sc.prim_mat[prim]];
        if (rk.D == 1 || !mat_is_delta(m)) {
          ci = (size_t)i * S + k;
          const uint32_t slot = q.slot[cur][i];
          const uint32_t pixel = act[slot / (uint32_t)nb];
          const uint32_t n0 = q.rngn[cur][i];
          const d3 hit_p = D3(q.h_geo[i], q.h_geo[cap + i], q.h_geo[2 * cap + i]);
          const d3 wo_world = D3(q.h_geo[3 * cap + i], q.h_geo[4 * cap + i], q.h_geo[5 * cap + i]);
          const d3 nrm = D3(q.h_geo[6 * cap + i], q.h_geo[7 * cap + i], q.h_geo[8 * cap + i]);
          d3 o2w[3]; make_coord_space(o2w, nrm);
          const d3 w_out = mat_tmul(o2w, wo_world);
          if (rk.hemi) {
            Rng g = rng_make(rk.seed, pixel, (uint32_t)(base_sample + slot % (uint32_t)nb), n0 + 2u * (uint32_t)k);
            d3 w_in = hemisphere_sample(g);
            wi_world = mat_mul(o2w, w_in);
            f3 f = bsdf_f(m, w_out, w_in);
            s_contrib[ci] = make_float4(f.r, f.g, f.b, (float)w_in.z);
            s_res[ci] = -1;
            dist_d = INFINITY; emit = true;
          } else {
            /* which light, which of its samples, and how many draws precede it */
            const int n_all = sc.n_lights + (sc.env ? 1 : 0);
            int li = 0, first = 0; uint32_t off = 0;
            bool is_env = false; int ns = 1, draws = 0;
            for (; li < n_all; ++li) {
              is_env = li >= sc.n_lights;
              const int ty = is_env ? -1 : sc.lights[li].type;
              ns = (!is_env && sc.lights[li].is_delta) ? 1 : rk.ns_area_light;
              draws = (is_env || ty == RRT_LIGHT_AREA || ty == RRT_LIGHT_HEMISPHERE) ? 2 : 0;
              if (k < first + ns) break;
              first += ns; off += (uint32_t)(ns * draws);
            }
            off += (uint32_t)((k - first) * draws);
            Rng g = rng_make(rk.seed, pixel, (uint32_t)(base_sample + slot % (uint32_t)nb), n0 + off);
            float dist, pdf;
            f3 Li = is_env ? env_sample_L(sc, wi_world, dist, pdf, g) : light_sample_L(sc.lights[li], hit_p, wi_world, dist, pdf, g);
            int verdict = -2;
            if (is_env || sc.lights[li].type != RRT_LIGHT_NULL) {
              d3 w_in = mat_tmul(o2w, wi_world);
              if (!(w_in.z < 0)) {
                f3 c = s_divf(s_mulf(s_mul(Li, bsdf_f(m, w_out, w_in)), (float)w_in.z), pdf);
                s_contrib[ci] = make_float4(c.r, c.g, c.b, 0.f);
                verdict = -1; dist_d = (double)dist; emit = true;
              }
            }
            s_res[ci] = verdict;
          }
          so = vadd(hit_p, smul(RRT_EPS_D, wi_world));
        }
      }
    }
    const unsigned mask = __ballot_sync(0xffffffffu, emit);
    if (mask) {
      int base = 0;
      if (lane_id() == 0) base = atomicAdd(sh_count, __popc(mask));
      base = __shfl_sync(0xffffffffu, base, 0);
      if (emit) {
        const int pos = base + __popc(mask & ((1u << lane_id()) - 1u));
        const size_t cs = cap * (size_t)S;
        sray[pos] = so.x; sray[cs + pos] = so.y; sray[2 * cs + pos] = so.z;
        sray[3 * cs + pos] = wi_world.x; sray[4 * cs + pos] = wi_world.y; sray[5 * cs + pos] = wi_world.z;
        sray[6 * cs + pos] = dist_d;
        s_cidx[pos] = (uint32_t)ci;
      }
    }
  }
}

/* ---------------------------------------------------------------- shadow rays in the order of their origin's cell (RRT_OPT_SORT_SHADOW).
 * The shadow rays of the second and later vertices of a path start anywhere in the scene, in the order the paths happen to sit in the queue:
 * the 32 rays a warp of the trace kernel holds have nothing in common, each needs the tree in a different place and for a different number of
 * visits, and a warp's trip through the traversal lasts as long as its slowest lane.  Handing the same rays out cell by cell — a 16 x 16 x 16
 * grid over the scene box, cells in Morton order, 12 cm at Cornell-box scale — makes the launch 10-12 % shorter (tools/sort_probe.py: finer
 * cells and the direction octant add nothing).  The rays stay where k_shade_direct put them; a counting sort over the cell numbers builds the
 * launch order `perm`, and the trace kernel's IO policy reads ray perm[i] where it read ray i.  Which lane traces a ray changes, no result does.
 * Three small kernels per launch, the queue's length read on the device like everywhere else: cell numbers + per-block histogram in shared
 * memory, a one-block scan of the 4096 counts, and the scatter (one atomic per group of lanes of a warp that share a cell). */
#define RRT_BIN_BITS 4
#define RRT_BINS (1 << (3 * RRT_BIN_BITS))
struct BinGrid { float lo[3], inv[3]; };   /* cell coordinate = (p - lo) * inv, clamped to the grid */

__device__ __forceinline__ uint32_t bin_spread(uint32_t v) {   /* bit k -> bit 3k */
  uint32_t r = 0;
#pragma unroll
  for (int k = 0; k < RRT_BIN_BITS; ++k) r |= ((v >> k) & 1u) << (3 * k);
  return r;
}
__device__ __forceinline__ uint32_t bin_of(const BinGrid& g, double x, double y, double z) {
  const int m = (1 << RRT_BIN_BITS) - 1;
  const int ix = min(max((int)(((float)x - g.lo[0]) * g.inv[0]), 0), m), iy = min(max((int)(((float)y - g.lo[1]) * g.inv[1]), 0), m),
            iz = min(max((int)(((float)z - g.lo[2]) * g.inv[2]), 0), m);   /* a NaN lands in cell 0 */
  return bin_spread((uint32_t)ix) | (bin_spread((uint32_t)iy) << 1) | (bin_spread((uint32_t)iz) << 2);
}
/* lanes of a warp that hold the same cell make ONE addition between them (rays reach the queue vertex by vertex and pixel by pixel: most of a warp
 * usually shares its cell, and 32 atomics on one word are 32 turns); returns the value before the addition + this lane's rank among its peers */
__device__ __forceinline__ uint32_t bin_add_grouped(uint32_t* counters, uint32_t b, unsigned act) {
  const unsigned peers = __match_any_sync(act, b);
  const int leader = __ffs(peers) - 1;
  uint32_t pos = 0;
  if (lane_id() == leader) pos = atomicAdd(&counters[b], (uint32_t)__popc(peers));
  pos = __shfl_sync(peers, pos, leader);
  return pos + (uint32_t)__popc(peers & ((1u << lane_id()) - 1u));
}
__global__ void __launch_bounds__(256) k_bin_count(const double* __restrict__ sray, size_t cs, const int* __restrict__ count_ptr, const BinGrid g,
                                                   uint32_t* __restrict__ hist, uint16_t* __restrict__ bin) {
  __shared__ uint32_t h[RRT_BINS];
  for (int t = threadIdx.x; t < RRT_BINS; t += blockDim.x) h[t] = 0;
  __syncthreads();
  const int n = *count_ptr;
  const int span = (n + 31) & ~31;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < span; i += gridDim.x * blockDim.x) {
    const bool ok = i < n;
    const unsigned act = __ballot_sync(0xffffffffu, ok);
    if (ok) {
      const uint32_t b = bin_of(g, sray[i], sray[cs + i], sray[2 * cs + i]);
      bin[i] = (uint16_t)b;
      bin_add_grouped(h, b, act);
    }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < RRT_BINS; t += blockDim.x)
    if (h[t]) atomicAdd(&hist[t], h[t]);
}
/* one block of 1024 threads: exclusive scan of the counts into `base`, and the counts back to zero for the next launch */
__global__ void __launch_bounds__(1024) k_bin_scan(uint32_t* __restrict__ hist, uint32_t* __restrict__ base) {
  constexpr int PER = RRT_BINS / 1024;
  static_assert(RRT_BINS % 1024 == 0, "k_bin_scan: 1024 threads, a whole number of cells each");
  __shared__ uint32_t wsum[32];
  uint32_t v[PER], sum = 0;
#pragma unroll
  for (int k = 0; k < PER; ++k) { v[k] = hist[threadIdx.x * PER + k]; hist[threadIdx.x * PER + k] = 0; sum += v[k]; }
  uint32_t inc = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, inc, o); if (lane_id() >= o) inc += y; }
  if (lane_id() == 31) wsum[threadIdx.x >> 5] = inc;
  __syncthreads();
  if (threadIdx.x < 32) {
    uint32_t w = wsum[threadIdx.x], wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, wi, o); if (lane_id() >= o) wi += y; }
    wsum[threadIdx.x] = wi - w;
  }
  __syncthreads();
  uint32_t ex = wsum[threadIdx.x >> 5] + inc - sum;
#pragma unroll
  for (int k = 0; k < PER; ++k) { base[threadIdx.x * PER + k] = ex; ex += v[k]; }
}
/* scatter, a tile of 4096 consecutive rays per block and turn: ranks within the tile from a histogram in shared memory, then ONE reservation per
 * (tile, cell) in the global table — the rays of a tile share a handful of cells when the queue is in pixel order, and an atomic per group of lanes
 * on those few words would run one after the other across the whole grid */
#define RRT_BIN_RPT 16
__global__ void __launch_bounds__(256) k_bin_scatter(const uint16_t* __restrict__ bin, const int* __restrict__ count_ptr, uint32_t* __restrict__ base,
                                                     uint32_t* __restrict__ perm) {
  __shared__ uint32_t h[RRT_BINS];
  const int n = *count_ptr;
  constexpr int TILE = 256 * RRT_BIN_RPT;
  for (int t0 = blockIdx.x * TILE; t0 < n; t0 += gridDim.x * TILE) {
    for (int t = threadIdx.x; t < RRT_BINS; t += 256) h[t] = 0;
    __syncthreads();
    uint32_t br[RRT_BIN_RPT];   /* cell << 16 | rank within (tile, cell) */
#pragma unroll
    for (int k = 0; k < RRT_BIN_RPT; ++k) {
      const int i = t0 + k * 256 + (int)threadIdx.x;
      const bool ok = i < n;
      const unsigned act = __ballot_sync(0xffffffffu, ok);
      br[k] = 0;
      if (ok) {
        const uint32_t b = bin[i];
        br[k] = (b << 16) | bin_add_grouped(h, b, act);
      }
    }
    __syncthreads();
    for (int t = threadIdx.x; t < RRT_BINS; t += 256) {
      const uint32_t c = h[t];
      if (c) h[t] = atomicAdd(&base[t], c);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < RRT_BIN_RPT; ++k) {
      const int i = t0 + k * 256 + (int)threadIdx.x;
      if (i < n) perm[h[br[k] >> 16] + (br[k] & 0xffffu)] = (uint32_t)i;
    }
    __syncthreads();
  }
}

/* ---------------------------------------------------------------- accumulate: fold the shadow-ray verdicts of vertex
 * `it` into each path's radiance, in the reference's sample order */
__global__ void __launch_bounds__(256) k_accumulate(const __grid_constant__ DevScene sc, const __grid_constant__ RenderK rk,
                                                    const __grid_constant__ Queues q, int it) {
  const int sp = it & 1;
  const size_t cap = (size_t)rk.cap;
  const int count = q.ctl[it];
  const uint8_t* const vflag = q.vflag[sp]; const float* const dbeta = q.dbeta[sp]; const float* const emit = q.emit[sp];
  const int* const s_res = q.s_res[sp]; const float4* const s_contrib = q.s_contrib[sp];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
    const uint8_t vf = vflag[i];
    if (!vf) continue;
    const uint32_t slot = q.aslot[sp][i];
    if (vf & 2) {   /* emission of a vertex reached through a delta BSDF (reference src/part1_code.cpp:94-95): before this vertex's direct light */
      q.Lsample[slot] += emit[i]; q.Lsample[cap + slot] += emit[cap + i]; q.Lsample[2 * cap + slot] += emit[2 * cap + i];
    }
    if (!(vf & 1)) continue;
    f3 L = F3(0, 0, 0);
    for (int k = 0; k < rk.S; ++k) {
      const size_t ci = (size_t)i * rk.S + k;
      const int r = s_res[ci];
      if (rk.hemi) {
        if (r >= 0) {
          float4 c = s_contrib[ci];
          f3 Le = mat_emission(sc.mats[sc.prim_mat[r]]);
          L = s_add(L, s_mulf(s_mul(Le, F3(c.x, c.y, c.z)), c.w));
        }
      } else if (r == -1) {
        float4 c = s_contrib[ci];
        L = s_add(L, F3(c.x, c.y, c.z));
      }
    }
    if (rk.hemi) L = s_divf(s_mulf(s_mulf(L, 2.f), (float)RRT_PI), (float)rk.total);
    else L = s_divf(L, (float)rk.total);
    const f3 beta = F3(dbeta[i], dbeta[cap + i], dbeta[2 * cap + i]);
    f3 a = s_mul(L, beta);
    q.Lsample[slot] += a.r; q.Lsample[cap + slot] += a.g; q.Lsample[2 * cap + slot] += a.b;
  }
}

/* ---------------------------------------------------------------- resolve: add a round's samples to their pixel in
 * sample order and apply the adaptive stop rule (reference src/part1_code.cpp:141-161) */
__global__ void __launch_bounds__(256) k_resolve(const __grid_constant__ RenderK rk, const __grid_constant__ Queues q, const __grid_constant__ PixelAcc px,
                                                 const uint32_t* __restrict__ act, int npix, int nb) {
  const size_t cap = (size_t)rk.cap, np = (size_t)rk.W * rk.H;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += gridDim.x * blockDim.x) {
    const uint32_t pixel = act[p];
    f3 ret = F3(px.sum[pixel], px.sum[np + pixel], px.sum[2 * np + pixel]);
    double s1 = px.s1[pixel], s2 = px.s2[pixel];
    int n = px.n[pixel];
    for (int s = 0; s < nb; ++s) {
      const size_t slot = (size_t)p * nb + s;
      f3 L = F3(q.Lsample[slot], q.Lsample[cap + slot], q.Lsample[2 * cap + slot]);
      ret = s_add(ret, L);
      double illum = s_illum(L);
      s1 += illum; s2 += illum * illum;
    }
    n += nb;
    if (n % rk.batch == 0) {
      double avg = s1 / n, sd = sqrt((s2 - avg * s1) / (n - 1));
      if (1.96 * sd / sqrt((double)n) <= rk.tol * avg) px.done[pixel] = 1;
    }
    if (n >= rk.ns_aa) px.done[pixel] = 1;
    px.sum[pixel] = ret.r; px.sum[np + pixel] = ret.g; px.sum[2 * np + pixel] = ret.b;
    px.s1[pixel] = s1; px.s2[pixel] = s2; px.n[pixel] = n;
  }
}

/* mean radiance and sample count per pixel (reference src/part1_code.cpp:161-162) */
__global__ void __launch_bounds__(256) k_finalize(const __grid_constant__ PixelAcc px, size_t npix, float* __restrict__ rgb, int* __restrict__ counts,
                                                  float4* __restrict__ rgbn, Counters* gcnt) {
  unsigned long long paths = 0;
  for (size_t p = blockIdx.x * (size_t)blockDim.x + threadIdx.x; p < npix; p += (size_t)gridDim.x * blockDim.x) {
    int n = px.n[p];
    paths += (unsigned long long)n;
    f3 s = F3(px.sum[p], px.sum[npix + p], px.sum[2 * npix + p]);
    f3 m = n > 0 ? s_divf(s, (float)n) : F3(0, 0, 0);
    if (rgb) { rgb[3 * p] = m.r; rgb[3 * p + 1] = m.g; rgb[3 * p + 2] = m.b; }
    if (counts) counts[p] = n;
    if (rgbn) rgbn[p] = make_float4(m.r, m.g, m.b, (float)n);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) paths += __shfl_xor_sync(0xffffffffu, paths, o);
  if (gcnt && lane_id() == 0 && paths) atomicAdd(&gcnt->paths, paths);   /* gcnt == nullptr: rrt_snapshot */
}

/* ---------------------------------------------------------------- stable compaction (ballot + prefix), three passes:
 * per-block counts, one-block exclusive scan of the counts, scatter.  keep(i) decides. */
#define RRT_CBLOCK 1024
/* Tile (tx, ty) belongs to rank (tx + ty * stride) % world, stride = rrt_shard_stride(world): the smallest integer >= round(sqrt(world))
 * that is coprime with world (8 -> 3, 4 -> 3, 2 -> 1, 6 -> 5).  Every run of `world` consecutive tiles of a row AND of a column then holds
 * every rank once, whatever the frame width.  (Round 1 used ty * tiles-per-row: with 1920 / 16 = 120 tiles per row and 8 ranks that is
 * tx % 8 — vertical stripes, 3 or 4 of the 30 tile columns the Cornell box covers per rank: 476-590 ms per rank on C4, SCALE 5.9x at N = 8.) */
__host__ __device__ inline int rrt_shard_stride(int world) {
  int s = 1;
  while ((s + 1) * (s + 1) <= world + s) ++s;     /* s = round(sqrt(world)) for the small counts that matter; any s >= 1 works */
  for (;; ++s) {
    int a = s, b = world;
    while (b) { int t = a % b; a = b; b = t; }
    if (a == 1) return s;
  }
}
struct KeepOwned {  // pixel belongs to this shard and to the render cell
  int W, x0, y0, x1, y1, tile, ntx /* = shard stride */, rank, world;
  __device__ bool operator()(uint32_t p, uint32_t) const {
    int x = (int)(p % (uint32_t)W), y = (int)(p / (uint32_t)W);
    if (x < x0 || x >= x1 || y < y0 || y >= y1) return false;
    return ((x / tile) + (y / tile) * ntx) % world == rank;
  }
};
struct KeepActive {  // pixel has not reached its stop rule
  const uint8_t* done;
  __device__ bool operator()(uint32_t, uint32_t v) const { return !done[v]; }
};

template <class Keep>
__global__ void __launch_bounds__(RRT_CBLOCK) k_compact_count(Keep keep, const uint32_t* __restrict__ in, uint32_t n, uint32_t* __restrict__ block_counts) {
  __shared__ uint32_t wsum[32];
  uint32_t i = blockIdx.x * RRT_CBLOCK + threadIdx.x;
  bool k = i < n && keep(i, in ? in[i] : i);
  unsigned m = __ballot_sync(0xffffffffu, k);
  if (lane_id() == 0) wsum[threadIdx.x >> 5] = __popc(m);
  __syncthreads();
  if (threadIdx.x < 32) {
    uint32_t v = wsum[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = v;
  }
}
__global__ void __launch_bounds__(1024) k_compact_scan(uint32_t* __restrict__ block_counts, uint32_t nblocks, uint32_t* __restrict__ total) {
  __shared__ uint32_t wsum[32];
  __shared__ uint32_t carry;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < nblocks; base += 1024) {
    uint32_t i = base + threadIdx.x;
    uint32_t v = i < nblocks ? block_counts[i] : 0, x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane_id() >= o) x += y; }
    if (lane_id() == 31) wsum[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      uint32_t w = wsum[threadIdx.x], z = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, z, o); if (lane_id() >= o) z += y; }
      wsum[threadIdx.x] = z - w;
    }
    __syncthreads();
    uint32_t excl = carry + wsum[threadIdx.x >> 5] + x - v;
    if (i < nblocks) block_counts[i] = excl;
    __syncthreads();
    if (threadIdx.x == 1023) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}
template <class Keep>
__global__ void __launch_bounds__(RRT_CBLOCK) k_compact_scatter(Keep keep, const uint32_t* __restrict__ in, uint32_t n, const uint32_t* __restrict__ block_offsets,
                                                                uint32_t* __restrict__ out) {
  __shared__ uint32_t wsum[32];
  uint32_t i = blockIdx.x * RRT_CBLOCK + threadIdx.x;
  uint32_t v = i < n ? (in ? in[i] : i) : 0;
  bool k = i < n && keep(i, v);
  unsigned m = __ballot_sync(0xffffffffu, k);
  if (lane_id() == 0) wsum[threadIdx.x >> 5] = __popc(m);
  __syncthreads();
  if (threadIdx.x < 32) {
    uint32_t w = wsum[threadIdx.x], z = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, z, o); if (lane_id() >= o) z += y; }
    wsum[threadIdx.x] = z - w;
  }
  __syncthreads();
  if (k) out[block_offsets[blockIdx.x] + wsum[threadIdx.x >> 5] + __popc(m & ((1u << lane_id()) - 1u))] = v;
}

/* ---------------------------------------------------------------- parity hooks (one thread per item) */
__global__ void k_hook_geodesic(const __grid_constant__ DevHole bh, long long n, const double* __restrict__ o, const double* __restrict__ d, int nsteps,
                                double* __restrict__ seg) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    d3 ro = D3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), rd = D3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
    double len = 0.0;
    for (int j = 0; j < nsteps; ++j) {
      next_micro_ray(bh, ro, rd, len);
      double* s = seg + ((size_t)i * nsteps + j) * 8;
      s[0] = ro.x; s[1] = ro.y; s[2] = ro.z; s[3] = rd.x; s[4] = rd.y; s[5] = rd.z; s[6] = len;
      s[7] = horizon_hit(bh, ro, rd, len) ? 1.0 : 0.0;
    }
  }
}

/* the same hook through the FP32 stepper (RRT_OPT_STEPPER = 1) */
__global__ void k_hook_geodesic_fp32(const __grid_constant__ DevHole bh, long long n, const double* __restrict__ o, const double* __restrict__ d, int nsteps,
                                     double* __restrict__ seg) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    d3 ro = D3(o[3 * i], o[3 * i + 1], o[3 * i + 2]), rd = D3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
    FastRay fr; fast_init(bh, ro, rd, fr);
    for (int j = 0; j < nsteps; ++j) {
      float ox, oy, oz, dx, dy, dz, len;
      const bool captured = fast_step(bh, fr, ox, oy, oz, dx, dy, dz, len);
      double* s = seg + ((size_t)i * nsteps + j) * 8;
      s[0] = ox; s[1] = oy; s[2] = oz; s[3] = dx; s[4] = dy; s[5] = dz; s[6] = len;
      s[7] = captured ? 1.0 : 0.0;
    }
  }
}

__global__ void k_hook_camera(const __grid_constant__ DevCamera cam, long long n, const double* __restrict__ xy, double* __restrict__ o, double* __restrict__ d) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    d3 ro, rd; generate_ray(cam, xy[2 * i], xy[2 * i + 1], ro, rd);
    o[3 * i] = ro.x; o[3 * i + 1] = ro.y; o[3 * i + 2] = ro.z; d[3 * i] = rd.x; d[3 * i + 1] = rd.y; d[3 * i + 2] = rd.z;
  }
}

__global__ void k_hook_camera_lens(const __grid_constant__ DevCamera cam, long long n, const double* __restrict__ xy, const double* __restrict__ lens,
                                   double* __restrict__ o, double* __restrict__ d) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    d3 ro, rd; generate_ray_thin_lens(cam, xy[2 * i], xy[2 * i + 1], lens[2 * i], lens[2 * i + 1], ro, rd);
    o[3 * i] = ro.x; o[3 * i + 1] = ro.y; o[3 * i + 2] = ro.z; d[3 * i] = rd.x; d[3 * i + 1] = rd.y; d[3 * i + 2] = rd.z;
  }
}

__global__ void k_hook_bsdf(const rrt_material* __restrict__ mats, int mat, long long n, const double* __restrict__ wo, const double* __restrict__ wi_in,
                            const double* __restrict__ u, float* __restrict__ f, double* __restrict__ wi, float* __restrict__ pdf) {
  const rrt_material m = mats[mat];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    d3 w = D3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]);
    f3 s;
    if (u) {  /* sample_f */
      Rng g = rng_make(0, 0, 0, 0); g.fixed = u + 4 * i;
      d3 r = D3(0, 0, 0); float p = 0;
      s = bsdf_sample_f(m, w, r, p, g);
      wi[3 * i] = r.x; wi[3 * i + 1] = r.y; wi[3 * i + 2] = r.z; pdf[i] = p;
    } else s = bsdf_f(m, w, D3(wi_in[3 * i], wi_in[3 * i + 1], wi_in[3 * i + 2]));
    f[3 * i] = s.r; f[3 * i + 1] = s.g; f[3 * i + 2] = s.b;
  }
}

/* parity hooks for the environment light: sample_L with caller-supplied uniforms, and sample_dir */
__global__ void k_hook_env_sample(const __grid_constant__ DevScene sc, long long n, const double* __restrict__ u, float* __restrict__ rad,
                                  double* __restrict__ wi, float* __restrict__ dist, float* __restrict__ pdf) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    Rng g = rng_make(0, 0, 0, 0); g.fixed = u + 2 * i;
    d3 w; float ds = 0, pd = 0;
    f3 s = env_sample_L(sc, w, ds, pd, g);
    rad[3 * i] = s.r; rad[3 * i + 1] = s.g; rad[3 * i + 2] = s.b;
    wi[3 * i] = w.x; wi[3 * i + 1] = w.y; wi[3 * i + 2] = w.z; dist[i] = ds; pdf[i] = pd;
  }
}
__global__ void k_hook_env_lookup(const __grid_constant__ DevScene sc, long long n, const double* __restrict__ d, float* __restrict__ rad) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    f3 s = env_sample_dir(sc, D3(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
    rad[3 * i] = s.r; rad[3 * i + 1] = s.g; rad[3 * i + 2] = s.b;
  }
}

__global__ void k_hook_light(const rrt_light* __restrict__ lights, int light, long long n, const double* __restrict__ p, const double* __restrict__ u,
                             float* __restrict__ rad, double* __restrict__ wi, float* __restrict__ dist, float* __restrict__ pdf) {
  const rrt_light l = lights[light];
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    Rng g = rng_make(0, 0, 0, 0); g.fixed = u + 2 * i;
    d3 w; float ds = 0, pd = 0;
    f3 s = light_sample_L(l, D3(p[3 * i], p[3 * i + 1], p[3 * i + 2]), w, ds, pd, g);
    rad[3 * i] = s.r; rad[3 * i + 1] = s.g; rad[3 * i + 2] = s.b;
    wi[3 * i] = w.x; wi[3 * i + 1] = w.y; wi[3 * i + 2] = w.z; dist[i] = ds; pdf[i] = pd;
  }
}

__global__ void k_hook_philox(unsigned long long seed, long long n, const uint32_t* __restrict__ ctr, uint32_t* __restrict__ out) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    uint32_t c0 = ctr[4 * i], c1 = ctr[4 * i + 1], c2 = ctr[4 * i + 2], c3 = ctr[4 * i + 3];
    philox4x32_10(c0, c1, c2, c3, (uint32_t)seed, (uint32_t)(seed >> 32));
    out[4 * i] = c0; out[4 * i + 1] = c1; out[4 * i + 2] = c2; out[4 * i + 3] = c3;
  }
}

/* HDRImageBuffer::toColor + ImageBuffer::update_pixel (reference src/image.h:183-198,53-62): gamma 2.2, level 1,
 * upper clamp only, truncation to 8 bits, alpha 0xFF */
__global__ void k_tonemap(long long n, const float* __restrict__ hdr, uint32_t* __restrict__ rgba) {
  const float one_over_gamma = 1.0f / 2.2f;
  const float exposure = (float)sqrt(pow(2.0, (double)1.0f));
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    uint32_t p = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      float c = powf(hdr[3 * i + k] * exposure, one_over_gamma);
      float v = c < 1.f ? c : 1.f;
      if (!(v >= 0.f)) v = 0.f;
      p |= ((uint32_t)(v * 255)) << (8 * k);
    }
    rgba[i] = p | 0xFF000000u;
  }
}
