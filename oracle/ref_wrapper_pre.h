// oracle/ref_wrapper_pre.h + ref_wrapper_post.h -- TEST INFRASTRUCTURE (parity oracle + CPU baseline), not product code.
//
// Compiles the reference translation unit (/root/reference/jpeg.cpp and its headers huffman / file /
// types / logger / utils, read from where they lie; not one line of its decode path is touched) into a headless command-line tool, oracle/_ref/ref_decode.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// execute it.  Nothing under jpeg_decoder_b200/ links or calls it.
//
// Accommodations:
//   * The X11 window (include/display.hpp) is NOT built: no X headers, no stand-in for them.  The build
//     recipe (oracle/Makefile) streams jpeg.cpp to the compiler with its one line
//     `#include "include/display.hpp"` (jpeg.cpp:11) swapped for `#include "include/types.hpp"` (the
//     only thing of display.hpp the decode path needs) and ref_nodisplay.h, whose displayImage()
//     returns false -- "Image Display Failed", jpeg.cpp:819-823.  Nothing is copied into the repo.
//   * -U_FORTIFY_SOURCE: utils.hpp:18-23 sprintf()s 3 bytes into new unsigned char[2].
//   * `#define class struct` so the private stages of `Image` (jpeg.cpp:19-789) can be driven
//     one at a time and the MCU array dumped between decodeHuffman() (jpeg.cpp:405-446) and
//     dequantize()/inverseDCT()/YCbCrToRGB() (jpeg.cpp:563-590, 735-753, 544-561).
//   * No -march=native / -mfma / -ffast-math: FMA contraction changes ~3 ppm of samples.
//
// Modes
//   ref_decode dump  <in.jpg> <out.coef> <out.rgb>
//        out.coef = raw MCU[bwR*bhR] after decodeHuffman(): int32 [pos][3][64]   (types.hpp:32-50)
//        out.rgb  = uint8 [H][W][3] gathered as display.hpp:19-34 does
//        prints one JSON line with the geometry and per-stage milliseconds
//   ref_decode full  <in.jpg> <out.rgb>
//        runs the reference's own readJPEG() (jpeg.cpp:826-907) end to end, then gathers RGB
//   ref_decode bench <reps> <file>...
//        times ctor + readJPEG() (parse + unstuff + Huffman + dequant + IDCT + colour) per file;
//        prints one JSON line {"images":..,"pixels":..,"seconds":..}
#include <iostream>
#include <fstream>
#include <string>
#include <cstring>
#include <math.h>
#include <map>
#include <vector>
#include <chrono>
#include <thread>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include <sstream>

#define class struct
#define main ref_main
// the reference TU follows on the compiler's input (oracle/Makefile), then ref_wrapper_post.h
