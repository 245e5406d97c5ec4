// b200jpeg_cli -- the reference's main() (jpeg.cpp:916-929) over the B200 path:
//   ./b200jpeg <image_path> [more images ...] [--out-dir DIR] [--bmp] [--spec]
// Reads each JPEG, decodes it on GPU 0 and writes <name>.ppm (or .bmp) next to --out-dir instead
// of opening an X11 window (display.hpp:49-120 has no headless equivalent).  Exit code 1 on a usage
// error or an undecodable file, like the reference.
#include <iostream>
#include <string>
#include <vector>

#include "../include/b200jpeg.hpp"

int main(int argc, char **argv) {
  if (argc < 2) {
    std::cerr << "Usage: " << argv[0] << " <image_path> [...] [--out-dir DIR] [--bmp] [--spec]" << std::endl;
    return 1;
  }
  std::string out_dir = ".";
  bool bmp = false, spec = false;
  std::vector<std::string> paths;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    if (a == "--out-dir" && i + 1 < argc) out_dir = argv[++i];
    else if (a == "--bmp") bmp = true;
    else if (a == "--spec") spec = true;
    else paths.push_back(a);
  }
  try {
    std::vector<b200jpeg::Image> images;
    images.reserve(paths.size());
    for (const std::string &p : paths) {
      images.emplace_back(p);
      if (spec) images.back().parse_flags = B200JPEG_PARSE_SPEC;
      images.back().readHeaders();
    }
    std::vector<b200jpeg::Image *> ptrs;
    for (b200jpeg::Image &im : images) ptrs.push_back(&im);
    const std::vector<int> st = b200jpeg::Image::decode_batch(ptrs);
    int bad = 0;
    for (size_t i = 0; i < images.size(); ++i) {
      std::string base = paths[i].substr(paths[i].find_last_of('/') + 1);
      base = base.substr(0, base.find_last_of('.'));
      if (st[i]) { std::cerr << paths[i] << ": corrupt scan, status bits " << st[i] << std::endl; ++bad; continue; }
      const std::string out = out_dir + "/" + base + (bmp ? ".bmp" : ".ppm");
      if (bmp) images[i].saveToBMP(out); else images[i].savePPM(out);
      std::cout << paths[i] << ": " << images[i].image_width << "x" << images[i].image_height << " -> " << out << std::endl;
    }
    return bad ? 1 : 0;
  } catch (const b200jpeg::Error &e) {
    std::cerr << "ERROR: " << e.what() << std::endl;
    return 1;
  }
}
