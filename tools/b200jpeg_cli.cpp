// b200jpeg_cli -- the reference's main() (jpeg.cpp:916-929) over the B200 path:
//   ./b200jpeg <image_path> [more images ...] [--out-dir DIR] [--bmp] [--spec] [--gpus N] [--no-write]
// Reads the JPEGs (all of them at once: one pinned arena, host threads), decodes them on N GPUs (contiguous ranges of
// the files, one context and host thread per GPU) and writes <name>.ppm (or .bmp) into --out-dir instead of opening an
// X11 window (display.hpp:49-120; unavailable on this box).  Exit code 1 on a usage error or an undecodable file, like
// the reference.  --spec: the superset of what the reference accepts (grey, COM segments, any component ids, 16-bit DQT)
// INCLUDING progressive files, which the reference refuses (jpeg.cpp:69-73).
#include <chrono>
#include <iostream>
#include <string>
#include <vector>

#include "../include/b200jpeg.hpp"

int main(int argc, char **argv) {
  if (argc < 2) {
    std::cerr << "Usage: " << argv[0] << " <image_path> [...] [--out-dir DIR] [--bmp] [--spec] [--gpus N] [--no-write] [--repeat K]" << std::endl;
    return 1;
  }
  std::string out_dir = ".";
  bool bmp = false, spec = false, write = true;
  int gpus = 1, repeat = 1;
  std::vector<std::string> paths;
  for (int i = 1; i < argc; ++i) {
    const std::string a = argv[i];
    if (a == "--out-dir" && i + 1 < argc) out_dir = argv[++i];
    else if (a == "--gpus" && i + 1 < argc) gpus = std::atoi(argv[++i]);
    else if (a == "--bmp") bmp = true;
    else if (a == "--spec") spec = true;
    else if (a == "--no-write") write = false;
    else if (a == "--repeat" && i + 1 < argc) repeat = std::atoi(argv[++i]);   // measurements: decode the same files again (contexts and page cache warm)
    else paths.push_back(a);
  }
  if (gpus < 1 || paths.empty()) { std::cerr << "ERROR: nothing to do" << std::endl; return 1; }
  try {
    std::vector<int> devices;
    for (int k = 0; k < gpus; ++k) devices.push_back(k);
    b200jpeg::Batch b;                               // reused over the repetitions: the pinned arenas are allocated once
    for (int r = 1; r < repeat; ++r) {
      b200jpeg::decode_files_into(b, paths, devices, spec ? B200JPEG_PARSE_SPEC : B200JPEG_PARSE_REFERENCE);
      std::cout << "run " << r << ": read " << b.ms_read << " ms, parse " << b.ms_parse << " ms, alloc " << b.ms_alloc << " ms, decode " << b.ms_decode << " ms" << std::endl;
    }
    const auto t0 = std::chrono::steady_clock::now();
    b200jpeg::decode_files_into(b, paths, devices, spec ? B200JPEG_PARSE_SPEC : B200JPEG_PARSE_REFERENCE);
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    int bad = 0;
    double mp = 0;
    for (size_t i = 0; i < paths.size(); ++i) {
      std::string base = paths[i].substr(paths[i].find_last_of('/') + 1);
      base = base.substr(0, base.find_last_of('.'));
      if (b.parse_rc[i]) { std::cerr << paths[i] << ": not decodable, b200jpeg error " << b.parse_rc[i] << std::endl; ++bad; continue; }
      if (b.status[i]) { std::cerr << paths[i] << ": corrupt scan, status bits " << b.status[i] << std::endl; ++bad; continue; }
      mp += 1e-6 * b.width(i) * b.height(i);
      if (!write) continue;
      const std::string out = out_dir + "/" + base + (bmp ? ".bmp" : ".ppm");
      if (bmp) {                                     // through Image: jpeg.cpp:809-816 with the byte order put right
        b200jpeg::Image im(b.files.data() + b.file_off[i], b.file_len[i]);
        im.adopt(b.descs[i], b.rgb(i));
        im.saveToBMP(out);
      } else b.savePPM(i, out);
      std::cout << paths[i] << ": " << b.width(i) << "x" << b.height(i) << " -> " << out << std::endl;
    }
    std::cout << paths.size() - (size_t)bad << " of " << paths.size() << " files decoded on " << gpus << " GPU(s): " << mp << " MP in " << ms
              << " ms (files read " << b.ms_read << " + headers " << b.ms_parse << " + output allocation " << b.ms_alloc << " + decode host to host " << b.ms_decode
              << ")" << std::endl;
    return bad ? 1 : 0;
  } catch (const b200jpeg::Error &e) {
    std::cerr << "ERROR: " << e.what() << std::endl;
    return 1;
  }
}
