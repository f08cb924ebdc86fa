// Host-side check of the BVH8 builder + traversal core (the same __host__ __device__ code the CUDA kernels
// call) against a brute-force closest-hit search.  Test scaffolding: built and run by tests/test_bvh8_host.py.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#include "../../software-path-tracer_b200/csrc/bvh8.h"
#include "../../software-path-tracer_b200/csrc/bvh8_traverse.cuh"
#include "../../software-path-tracer_b200/csrc/kernels.h"

using namespace cudapt;

struct HostStack {
  uint2 e[64];
  int sp = 0, max_sp = 0;
  void push(uint2 v) {
    if (sp >= 64) {
      std::fprintf(stderr, "stack overflow\n");
      std::abort();
    }
    e[sp++] = v;
    if (sp > max_sp) max_sp = sp;
  }
  uint2 pop() { return e[--sp]; }
  bool empty() const { return sp == 0; }
};

static uint64_t g_state = 0x9E3779B97F4A7C15ull;
static float rnd() {
  g_state = g_state * 6364136223846793005ull + 1442695040888963407ull;
  return (float)((g_state >> 40) & 0xFFFFFF) / 16777216.0f;
}

struct Scene {
  std::vector<float> v;
  std::vector<uint32_t> idx;
};

static Scene soup(int n, float extent, float size) {
  Scene s;
  for (int i = 0; i < n; ++i) {
    const float cx = (rnd() - 0.5f) * extent, cy = (rnd() - 0.5f) * extent, cz = (rnd() - 0.5f) * extent + 20.f;
    for (int k = 0; k < 3; ++k) {
      s.v.push_back(cx + (rnd() - 0.5f) * size);
      s.v.push_back(cy + (rnd() - 0.5f) * size);
      s.v.push_back(cz + (rnd() - 0.5f) * size);
      s.idx.push_back((uint32_t)(3 * i + k));
    }
  }
  return s;
}
static Scene grid(int n) {  // height field + axis-aligned flat walls (zero-thickness boxes)
  Scene s;
  for (int z = 0; z <= n; ++z)
    for (int x = 0; x <= n; ++x) {
      s.v.push_back(-8.f + 16.f * x / n);
      s.v.push_back(-2.f + 0.7f * std::sin(0.9f * x) * std::cos(0.7f * z) + 0.2f * rnd());
      s.v.push_back(2.f + 16.f * z / n);
    }
  for (int z = 0; z < n; ++z)
    for (int x = 0; x < n; ++x) {
      const uint32_t a = z * (n + 1) + x, b = a + 1, c = a + n + 1, d = c + 1;
      s.idx.insert(s.idx.end(), {a, c, b, b, c, d});
    }
  const uint32_t base = (uint32_t)s.v.size() / 3;
  const float q[4][3] = {{-8, -3, 18}, {8, -3, 18}, {8, 5, 18}, {-8, 5, 18}};
  for (auto& p : q) s.v.insert(s.v.end(), {p[0], p[1], p[2]});
  s.idx.insert(s.idx.end(), {base, base + 1, base + 2, base, base + 2, base + 3});
  return s;
}

static int check(const char* name, Scene sc, int n_rays, bool from_inside, bool refit = false) {
  Bvh8 bvh;
  std::string err;
  const uint32_t nt = (uint32_t)sc.idx.size() / 3;
  if (!build_bvh8(sc.v.data(), (uint32_t)sc.v.size() / 3, sc.idx.data(), nt, nullptr, 4, &bvh, &err)) {
    std::printf("%s: build failed: %s\n", name, err.c_str());
    return 1;
  }
  if (refit) {  // move every vertex, refit (no rebuild) and validate the refitted tree exactly like a fresh one
    for (float& c : sc.v) c += (rnd() - 0.5f) * 0.6f;
    if (!refit_bvh8(sc.v.data(), (uint32_t)sc.v.size() / 3, sc.idx.data(), nt, &bvh, &err)) {
      std::printf("%s: refit failed: %s\n", name, err.c_str());
      return 1;
    }
    // the packed records must be exactly what a fresh build of the moved mesh would store
    Bvh8 fresh;
    if (!build_bvh8(sc.v.data(), (uint32_t)sc.v.size() / 3, sc.idx.data(), nt, nullptr, 4, &fresh, &err)) return 1;
    for (uint32_t i = 0; i < nt; ++i)
      if (std::memcmp(&fresh.tris[fresh.tri_of_prim[i]], &bvh.tris[bvh.tri_of_prim[i]], sizeof(PackedTri)) != 0) {
        std::printf("%s: refit triangle record %u differs from a fresh build\n", name, i);
        return 1;
      }
  }
  // structural checks
  std::vector<uint8_t> seen(nt, 0);
  for (const PackedTri& t : bvh.tris) seen[t.prim]++;
  for (uint32_t i = 0; i < nt; ++i)
    if (seen[i] != 1 || bvh.tris[bvh.tri_of_prim[i]].prim != i) {
      std::printf("%s: triangle %u appears %d times\n", name, i, seen[i]);
      return 1;
    }
  const uint4* nodes = reinterpret_cast<const uint4*>(bvh.nodes.data());
  const float4* tris = reinterpret_cast<const float4*>(bvh.tris.data());
  int bad = 0, hits = 0, max_sp = 0;
  double nodes_per_ray = 0, tris_per_ray = 0;
  for (int i = 0; i < n_rays; ++i) {
    f3 o, d;
    if (from_inside) {
      o = mk3((rnd() - 0.5f) * 10.f, (rnd() - 0.5f) * 10.f, 20.f + (rnd() - 0.5f) * 10.f);
      d = mk3(rnd() - 0.5f, rnd() - 0.5f, rnd() - 0.5f);
    } else {
      o = mk3(0, 0, 0);
      d = mk3((rnd() - 0.5f) * 1.6f, (rnd() - 0.5f) * 1.2f, 1.0f);
    }
    if (i % 17 == 0) d.x = 0.f;  // axis-parallel components
    if (i % 23 == 0) d.y = 0.f;
    d = normalize3(d);
    const float tnear = 1e-3f;
    // brute force with the same rule
    TraceHit want{std::numeric_limits<float>::infinity(), 0xFFFFFFFFu, 0xFFFFFFFFu};
    for (uint32_t k = 0; k < nt; ++k) {
      const PackedTri& T = bvh.tris[k];
      float t;
      if (!tri_test(mk3(T.v0[0], T.v0[1], T.v0[2]), mk3(T.e1[0], T.e1[1], T.e1[2]), mk3(T.e2[0], T.e2[1], T.e2[2]), o, d, t))
        continue;
      if (!(t >= tnear)) continue;
      if (t < want.t || (t == want.t && T.prim < want.prim)) want = TraceHit{t, k, T.prim};
    }
    HostStack st;
    TraceHit got;
    uint32_t nn = 0, ntr = 0;
    const RayPre r = prepare_ray(o, d);
    bvh8_trace<false, true>(nodes, tris, r, tnear, std::numeric_limits<float>::infinity(), st, got, nn, ntr);
    nodes_per_ray += nn;
    tris_per_ray += ntr;
    max_sp = st.max_sp > max_sp ? st.max_sp : max_sp;
    if (got.tri != want.tri || std::memcmp(&got.t, &want.t, 4) != 0) {
      if (bad < 5) std::printf("%s: ray %d mismatch: got tri %u t %.9g, want tri %u t %.9g\n", name, i, got.tri, got.t, want.tri, want.t);
      ++bad;
    }
    hits += want.tri != 0xFFFFFFFFu;
    // any-hit consistency on a shortened segment
    if (want.tri != 0xFFFFFFFFu) {
      HostStack s2;
      TraceHit occ;
      bvh8_trace<true, false>(nodes, tris, r, tnear, want.t * 1.01f, s2, occ, nn, ntr);
      if (occ.tri == 0xFFFFFFFFu) ++bad;
      HostStack s3;
      bvh8_trace<true, false>(nodes, tris, r, tnear, want.t * 0.99f, s3, occ, nn, ntr);
      if (occ.tri != 0xFFFFFFFFu && want.t * 0.99f > tnear) ++bad;
    }
  }
  if (std::getenv("BVH8_CHECK_STATS")) {  // slot usage of the tree (builder tuning)
    size_t internal = 0, leaf_slots = 0, leaf_tris = 0, leaf_only_nodes = 0;
    for (const Bvh8Node& N : bvh.nodes) {
      internal += (size_t)__builtin_popcount(N.imask);
      const uint32_t slots = (N.tri_valid | (N.tri_valid >> 8) | (N.tri_valid >> 16)) & 0xFFu;
      leaf_slots += (size_t)__builtin_popcount(slots);
      leaf_tris += (size_t)__builtin_popcount(N.tri_valid);
      if (!N.imask) ++leaf_only_nodes;
    }
    std::printf("%s: per node: %.2f internal children, %.2f leaf slots, %.2f triangles (%.2f per leaf slot); %.1f %% of the nodes have only leaves\n",
                name, (double)internal / bvh.nodes.size(), (double)leaf_slots / bvh.nodes.size(), (double)leaf_tris / bvh.nodes.size(),
                (double)leaf_tris / leaf_slots, 100.0 * leaf_only_nodes / bvh.nodes.size());
  }
  std::printf("%s: tris %u nodes %zu depth %u sah %.2f | rays %d hits %d mismatches %d | nodes/ray %.2f tris/ray %.2f max stack %d\n",
              name, nt, bvh.nodes.size(), bvh.max_depth, bvh.sah_cost, n_rays, hits, bad, nodes_per_ray / n_rays,
              tris_per_ray / n_rays, max_sp);
  return bad;
}

// FastDiv (kernels.h): the multiply-high division of the path-id decode against the machine's own division
static int check_fastdiv() {
  int bad = 0;
  uint64_t lcg = 0x9E3779B97F4A7C15ull;
  auto rnd = [&]() { lcg = lcg * 6364136223846793005ull + 1442695040888963407ull; return (uint32_t)(lcg >> 32); };
  std::vector<uint32_t> ds = {1u, 2u, 3u, 5u, 7u, 8u, 9u, 63u, 64u, 65u, 240u, 480u, 1920u, 3840u, 2073600u, 2088960u, 8294400u,
                              0x7FFFFFFFu, 0x80000000u, 0x80000001u, 0xFFFFFFFEu, 0xFFFFFFFFu};
  for (int i = 0; i < 400; ++i) ds.push_back(rnd() | 1u), ds.push_back(1u + rnd() % 5000u), ds.push_back(1u << (rnd() % 32u));
  for (uint32_t d : ds) {
    const cudapt::FastDiv f = cudapt::make_fastdiv(d);
    std::vector<uint32_t> xs = {0u, 1u, d - 1u, d, d + 1u, 2u * d - 1u, 2u * d, 0x7FFFFFFFu, 0x80000000u, 0xFFFFFFFEu, 0xFFFFFFFFu};
    for (int i = 0; i < 2000; ++i) xs.push_back(rnd());
    for (uint32_t k = 0; k < 64u; ++k) xs.push_back((uint32_t)((uint64_t)d * ((uint64_t)rnd() % (0xFFFFFFFFull / d + 1ull))) + (k % 3u) - 1u);
    for (uint32_t x : xs)
      if (cudapt::fd_div(f, x) != x / d) {
        if (++bad < 5) std::printf("fastdiv: %u / %u = %u, got %u\n", x, d, x / d, cudapt::fd_div(f, x));
      }
  }
  std::printf("fastdiv: %zu divisors, %s\n", ds.size(), bad ? "MISMATCH" : "exact");
  return bad;
}

int main(int argc, char** argv) {
  const int n_rays = argc > 1 ? std::atoi(argv[1]) : 3000;
  int bad = check_fastdiv();
  bad += check("soup-2k", soup(2000, 12.f, 1.5f), n_rays, true);
  bad += check("soup-sparse", soup(300, 30.f, 0.5f), n_rays, true);
  bad += check("grid-64", grid(64), n_rays, false);
  bad += check("tiny-1", soup(1, 1.f, 4.f), 200, true);
  bad += check("tiny-5", soup(5, 2.f, 4.f), 500, true);
  bad += check("refit-soup-2k", soup(2000, 12.f, 1.5f), n_rays, true, true);
  bad += check("refit-grid-64", grid(64), n_rays, false, true);
  bad += check("refit-tiny-1", soup(1, 1.f, 4.f), 200, true, true);
  std::printf(bad ? "FAILED (%d)\n" : "OK\n", bad);
  return bad ? 1 : 0;
}
