// vo_demo: the reference's src/demo.cpp loop (read raw I1_%06d.dat frames, process(), print the
// number of matches and inliers per frame, accumulate the pose as src/demo.cpp:222-236 does) on the C++ shim.
// Usage: vo_demo <dir> <width> <height> <n_frames> [f cu cv]
// Frames: <dir>/I1_%06d.dat (raw, as the reference), else <dir>/I1_%06d.png, else KITTI's <dir>/image_0/%06d.png or
// <dir>/%06d.png (right camera: I2_ / image_1); PNG files are decoded on the host (host/image_io.h), and width / height
// may then be given as 0.
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include <cstring>

#include "image_io.h"
#include "viso_mono.h"
#include "viso_stereo.h"

// one frame of camera `cam` (1 = left, 2 = right); width / height are checked against (or, if 0, taken from) the file
static bool load_frame(const std::string& dir, int cam, int32_t i, int32_t& width, int32_t& height, std::vector<uint8_t>& img) {
  char a[64], b[64], c[64], d[64];
  snprintf(a, sizeof(a), "/I%d_%06d.dat", cam, i);
  snprintf(b, sizeof(b), "/I%d_%06d.png", cam, i);
  snprintf(c, sizeof(c), "/image_%d/%06d.png", cam - 1, i);
  snprintf(d, sizeof(d), "/%06d.png", i);
  const char* names[4] = {a, b, c, cam == 1 ? d : nullptr};
  for (const char* nm : names) {
    if (!nm) continue;
    FILE* f = fopen((dir + nm).c_str(), "rb");
    if (!f) continue;
    fclose(f);
    int32_t w = width, h = height;
    std::string err;
    if (!vo_io::read_frame(dir + nm, img, &w, &h, &err)) { std::cerr << "ERROR: " << err << std::endl; return false; }
    if ((width > 0 && w != width) || (height > 0 && h != height)) {
      std::cerr << "ERROR: " << nm << " is " << w << "x" << h << ", expected " << width << "x" << height << std::endl;
      return false;
    }
    width = w; height = h;
    return true;
  }
  std::cerr << "ERROR: Couldn't read input files!" << std::endl;
  return false;
}

// stereo variant of the loop (src/demo.cpp reads I1_/I2_ pairs the same way): vo_demo <dir> w h n f cu cv --stereo <base>
static int run_stereo(const std::string& dir, int32_t width, int32_t height, int32_t n, double f, double cu, double cv, double base) {
  VisualOdometryStereo::parameters param;
  param.calib.f = f; param.calib.cu = cu; param.calib.cv = cv; param.base = base;
  VisualOdometryStereo viso(param);
  Matrix pose = Matrix::eye(4);
  std::vector<uint8_t> left, right;
  for (int32_t i = 0; i < n; i++) {
    if (!load_frame(dir, 1, i, width, height, left) || !load_frame(dir, 2, i, width, height, right)) return 1;
    int32_t dims[] = {width, height, width};
    const bool ok = viso.process(left.data(), right.data(), dims);
    if (viso.getMatcher()->lastStatus() != 0) return 2;
    std::cout << "Processing: Frame: " << i << ", Matches: " << viso.getNumberOfMatches();
    if (ok) {
      pose = pose * Matrix::inv(viso.getMotion());
      std::cout << ", Inliers: " << 100.0 * viso.getNumberOfInliers() / viso.getNumberOfMatches() << " %" << std::endl << pose << std::endl;
    } else {
      std::cout << " ... failed!" << std::endl;
    }
  }
  std::cout << "Demo complete! Exiting ..." << std::endl;
  return 0;
}

int main(int argc, char** argv) {
  if (argc < 5) {
    std::cerr << "Usage: vo_demo path/to/sequence width height n_frames" << std::endl;
    return 1;
  }
  const std::string dir = argv[1];
  int32_t width = atoi(argv[2]), height = atoi(argv[3]);
  const int32_t n = atoi(argv[4]);
  if (argc > 9 && !strcmp(argv[8], "--stereo"))
    return run_stereo(dir, width, height, n, atof(argv[5]), atof(argv[6]), atof(argv[7]), atof(argv[9]));
  VisualOdometryMono::parameters param;
  param.calib.f = argc > 5 ? atof(argv[5]) : 491.543832;   // calibration of the reference demo: src/demo.cpp:60-65
  param.calib.cu = argc > 6 ? atof(argv[6]) : 484.474328;
  param.calib.cv = argc > 7 ? atof(argv[7]) : 147.888234;
  param.height = 1.6;
  param.pitch = -0.08;
  VisualOdometryMono viso(param);
  Matrix pose = Matrix::eye(4);
  std::vector<uint8_t> img;
  for (int32_t i = 0; i < n; i++) {
    if (!load_frame(dir, 1, i, width, height, img)) return 1;
    int32_t dims[] = {width, height, width};
    const bool ok = viso.process(img.data(), dims);
    if (viso.getMatcher()->lastStatus() != 0) return 2;
    std::vector<Matcher::p_match> m = viso.getMatches();
    double su = 0, sv = 0;
    for (size_t k = 0; k < m.size(); k++) { su += m[k].u1c - m[k].u1p; sv += m[k].v1c - m[k].v1p; }
    std::cout << "Processing: Frame: " << i << ", Matches: " << m.size();
    if (!m.empty()) std::cout << ", mean flow: " << su / m.size() << " " << sv / m.size();
    if (ok) {  // src/demo.cpp:225-232
      pose = pose * Matrix::inv(viso.getMotion());
      std::cout << ", Inliers: " << 100.0 * viso.getNumberOfInliers() / viso.getNumberOfMatches() << " %" << std::endl << pose << std::endl;
    } else {
      std::cout << " ... failed!" << std::endl;
    }
  }
  std::cout << "Demo complete! Exiting ..." << std::endl;
  return 0;
}
