// Drives the C++ host shim (include/ydorb.hpp) the way the reference's Frame constructor drives its front-end
// (src/frame.cpp:60-106): two extractors on a stereo pair, computeStereoMatches, the key-point grid, a brute-force
// top-2 match and computeBoW.  Inputs come from files written by tests/test_gpu_cpp_shim.py, which compares the
// outputs written here with the CPU oracle.   usage: shim_check <in_dir> <out_dir>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "ydorb.hpp"

using namespace YDORBSLAM;

template <class T>
static std::vector<T> read_file(const std::string& path) {
  FILE* f = std::fopen(path.c_str(), "rb");
  if (!f) { std::fprintf(stderr, "cannot open %s\n", path.c_str()); std::exit(2); }
  std::fseek(f, 0, SEEK_END);
  const long sz = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  std::vector<T> v(sz / sizeof(T));
  if (sz > 0 && std::fread(v.data(), 1, sz, f) != (size_t)sz) std::exit(2);
  std::fclose(f);
  return v;
}

template <class T>
static void write_file(const std::string& path, const std::vector<T>& v) {
  FILE* f = std::fopen(path.c_str(), "wb");
  if (!f) std::exit(2);
  if (!v.empty()) std::fwrite(v.data(), sizeof(T), v.size(), f);
  std::fclose(f);
}

int main(int argc, char** argv) {
  if (argc < 3) return 2;
  const std::string in = std::string(argv[1]) + "/", out = std::string(argv[2]) + "/";
  try {
    const std::vector<int32_t> meta = read_file<int32_t>(in + "meta.bin");  // w, h, nFeatures, levelsup
    const std::vector<float> cam = read_file<float>(in + "cam.bin");        // bf, baseline
    const int w = meta[0], h = meta[1];
    const std::vector<uint8_t> imgL = read_file<uint8_t>(in + "left.bin"), imgR = read_file<uint8_t>(in + "right.bin");

    OrbExtractor left(meta[2], 1.2f, 8, 20, 7, w, h), right(meta[2], 1.2f, 8, 20, 7, w, h);
    std::vector<KeyPoint> kpsL, kpsR;
    std::vector<uint8_t> descL, descR;
    left.extractAndCompute(ImageView(imgL.data(), w, h), kpsL, descL);
    right.extractAndCompute(ImageView(imgR.data(), w, h), kpsR, descR);
    write_file(out + "kps_left.bin", kpsL);
    write_file(out + "desc_left.bin", descL);
    write_file(out + "kps_right.bin", kpsR);
    write_file(out + "desc_right.bin", descR);
    write_file(out + "scale_factors.bin", left.getScaleFactors());
    int pw = 0, ph = 0;
    write_file(out + "pyr3.bin", left.getImagePyramidLevel(3, &pw, &ph));
    std::vector<int32_t> pyr_dims = {pw, ph, left.getLevelsNum()};
    write_file(out + "pyr3_dims.bin", pyr_dims);

    OrbMatcher matcher(0.6f, true);
    std::vector<float> rightX, depth;
    computeStereoMatches(matcher, left, right, kpsL, descL, kpsR, descR, cam[0], cam[1], rightX, depth);
    write_file(out + "right_x.bin", rightX);
    write_file(out + "depth.bin", depth);

    FrameFeatures frame(matcher, kpsL, descL, rightX, 0.f, 0.f, (float)w, (float)h, left.getScaleFactors());
    const std::vector<float> q = read_file<float>(in + "area_queries.bin");  // x, y, r, minLevel, maxLevel per query
    std::vector<int32_t> area;
    for (size_t i = 0; i + 4 < q.size() + 1; i += 5) {
      const std::vector<int> idx = frame.getKeyPointsInArea(q[i], q[i + 1], q[i + 2], (int)q[i + 3], (int)q[i + 4]);
      area.push_back((int32_t)idx.size());
      area.insert(area.end(), idx.begin(), idx.end());
    }
    write_file(out + "area.bin", area);

    const std::vector<YdTop2> t2 = matcher.top2(descL.data(), (int)kpsL.size(), descR.data(), (int)kpsR.size());
    write_file(out + "top2.bin", t2);
    std::vector<uint8_t> acc;
    for (const YdTop2& t : t2) acc.push_back(matcher.accept(t) ? 1 : 0);
    write_file(out + "accept.bin", acc);
    std::vector<int32_t> one = {matcher.computeDescriptorsDistance(descL.data(), descR.data())};
    write_file(out + "dist0.bin", one);

    Vocabulary voc(matcher, in + "voc.dbow3");
    BowVector bow;
    FeatureVector fv;
    voc.transform(descL, bow, fv, meta[3]);
    write_file(out + "bow_ids.bin", bow.wordIds);
    write_file(out + "bow_values.bin", bow.values);
    write_file(out + "fv_nodes.bin", fv.nodeIds);
    write_file(out + "fv_start.bin", fv.nodeStart);
    write_file(out + "fv_items.bin", fv.items);

    // error behaviour: an empty image returns silently with no key points (src/orbExtractor.cpp:356-358) ...
    std::vector<KeyPoint> none;
    std::vector<uint8_t> noned;
    left.extractAndCompute(ImageView(), none, noned);
    if (!none.empty() || !noned.empty()) { std::fprintf(stderr, "empty image produced key points\n"); return 1; }
    // ... and a failing C-ABI call throws with the library's message instead of returning garbage
    bool threw = false;
    try {
      left.extractAndCompute(ImageView(imgL.data(), w * 4, h * 4), none, noned);  // larger than maxWidth x maxHeight
    } catch (const Error& e) {
      threw = e.code == YDORB_E_INVALID;
    }
    if (!threw) { std::fprintf(stderr, "oversized image did not throw\n"); return 1; }
    std::printf("shim ok: %zu + %zu key points, %zu words\n", kpsL.size(), kpsR.size(), bow.wordIds.size());
  } catch (const std::exception& e) {
    std::fprintf(stderr, "shim_check failed: %s\n", e.what());
    return 1;
  }
  return 0;
}
