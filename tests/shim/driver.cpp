// driver.cpp — TEST-ONLY: runs the C++ BundleAdjuster shim (include/bundle_adjuster.h) the way src/slam.cpp
// uses the reference class, on a map read from a flat binary file written by tests/test_shim.py.
//   shim_driver <in.bin> <out.bin> <start> <end> pack|optimize
// in:  i32 N, i32 M, f32 T_wc[N*16], i64 frame_off[N+1], u32 obs_pid[K], f32 obs_uv[K*4], f32 points[M*3],
//      f64 p_l[12], f64 p_r[12]
// out (pack):     i32 nc, i32 np, i64 nobs, f64 poses, f64 points, u32 ids, i32 cam, i32 pt, f32 uv
// out (optimize): i32 status, f64 final_cost, i32 iterations, f32 T_wc[N*16], f32 points[M*3]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <set>

#include "standins.h"
#include "../../include/bundle_adjuster.h"

using BundleAdjuster = sba::BundleAdjusterT<Frame, MapPoint>;   // the alias INTEGRATION.md adds to the reference

template <class T> static void rd(FILE* f, T* p, size_t n) { if (n && fread(p, sizeof(T), n, f) != n) { std::perror("read"); std::exit(2); } }
template <class T> static void wr(FILE* f, const T* p, size_t n) { if (n && fwrite(p, sizeof(T), n, f) != n) { std::perror("write"); std::exit(2); } }

int main(int argc, char** argv) {
  if (argc < 6) { std::fprintf(stderr, "usage: shim_driver in out start end pack|optimize\n"); return 2; }
  FILE* f = std::fopen(argv[1], "rb");
  if (!f) { std::perror(argv[1]); return 2; }
  int32_t N, M;
  rd(f, &N, 1); rd(f, &M, 1);
  std::vector<float> T((size_t)N * 16); rd(f, T.data(), T.size());
  std::vector<int64_t> off((size_t)N + 1); rd(f, off.data(), off.size());
  const size_t K = (size_t)off[N];
  std::vector<uint32_t> pid(K); rd(f, pid.data(), K);
  std::vector<float> uv(K * 4); rd(f, uv.data(), uv.size());
  std::vector<float> pts((size_t)M * 3); rd(f, pts.data(), pts.size());
  std::array<double, 12> pl, pr; rd(f, pl.data(), 12); rd(f, pr.data(), 12);
  std::fclose(f);

  // build the map the way VisualOdometer does (src/visual_odometer.cpp:335-444): points first seen in a frame are "first observed" there
  std::vector<MapPoint*> points; std::vector<Frame*> frames;
  for (int32_t i = 0; i < M; ++i) points.push_back(new MapPoint(pts[3 * (size_t)i], pts[3 * (size_t)i + 1], pts[3 * (size_t)i + 2]));
  std::set<uint32_t> seen;
  for (int32_t i = 0; i < N; ++i) {
    Mat4 P; std::memcpy(P.v, &T[(size_t)i * 16], 64);
    Frame* fr = new Frame(i ? frames[i - 1] : nullptr, P);
    for (int64_t k = off[i]; k < off[i + 1]; ++k) {
      fr->AddObservation(Observation(pid[k], uv[4 * k], uv[4 * k + 1], uv[4 * k + 2], uv[4 * k + 3], 1.0f));
      fr->AddMapPoints(points[pid[k]], seen.insert(pid[k]).second);
    }
    frames.push_back(fr);
  }
  sba::ReprojectionError::SetLeftProjection(pl);     // src/slam.cpp:176-209
  sba::ReprojectionError::SetRightProjection(pr);
  BundleAdjuster ba(frames, points);                 // src/slam.cpp:94
  ba.Options().max_time_s = 0.0;                     // parity runs: no wall-clock cap (src/params.h:40)
  const unsigned start = (unsigned)std::atoi(argv[3]), end = (unsigned)std::atoi(argv[4]);
  FILE* o = std::fopen(argv[2], "wb");
  if (!o) { std::perror(argv[2]); return 2; }
  try {
    if (!std::strcmp(argv[5], "pack")) {
      const int32_t np = ba.Pack(start, end), nc = (int32_t)(end - start);
      int64_t nobs = 0;
      for (unsigned i = start; i < end; ++i) nobs += (int64_t)frames[i]->Observations().size();
      wr(o, &nc, 1); wr(o, &np, 1); wr(o, &nobs, 1);
      wr(o, ba.PackedPoses().data(), (size_t)nc * 6); wr(o, ba.PackedPoints().data(), (size_t)np * 3);
      wr(o, ba.PackedPointIds().data(), (size_t)np); wr(o, ba.PackedObsCam().data(), (size_t)nobs);
      wr(o, ba.PackedObsPt().data(), (size_t)nobs); wr(o, ba.PackedObsUv().data(), (size_t)nobs * 4);
    } else {
      ba.Optimize(start, end);                       // src/slam.cpp:123,128,156
      const int32_t st = ba.Status(), it = ba.Summary().iterations;
      const double fc = ba.Summary().final_cost;
      wr(o, &st, 1); wr(o, &fc, 1); wr(o, &it, 1);
      for (int32_t i = 0; i < N; ++i) { const Mat4 P = frames[i]->GlobalPose(); wr(o, P.v, 16); }
      for (int32_t i = 0; i < M; ++i) { const auto p = points[i]->Position(); wr(o, p.data(), 3); }
    }
  } catch (const std::exception& e) {
    std::fprintf(stderr, "shim_driver: %s\n", e.what());
    std::fclose(o);
    return 3;
  }
  std::fclose(o);
  return 0;
}
