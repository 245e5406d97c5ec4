// oracle/ref_nodisplay.h -- TEST INFRASTRUCTURE.  Headless replacement of the one function jpeg.cpp uses
// from include/display.hpp (displayImage, display.hpp:49-120: the X11 window).  The display is
// unavailable on this box and is not on the scan-to-pixels path; Image::display() (jpeg.cpp:818-824)
// then reports "Image Display Failed", which is what the reference does when XOpenDisplay() fails.
#pragma once
static bool displayImage(MCU *, unsigned int, unsigned int, unsigned int, unsigned int, int) { return false; }
