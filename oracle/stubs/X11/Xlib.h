/* Headless stand-in for <X11/Xlib.h>, used ONLY to compile the untouched reference
 * (include/display.hpp:1) for the parity oracle in oracle/_ref.  No X library exists on
 * this box or the GPU box.  XOpenDisplay() returns NULL, so displayImage() takes its own
 * "Cannot open display" exit (display.hpp:56-61) and never reaches the other calls.
 * TEST INFRASTRUCTURE - never linked into the product library. */
#ifndef B200JPEG_ORACLE_XLIB_STUB_H
#define B200JPEG_ORACLE_XLIB_STUB_H
typedef struct _XDisplayStub Display;
typedef unsigned long Window;
typedef unsigned long Atom;
typedef unsigned long Colormap;
typedef struct _XGCStub *GC;
typedef struct { unsigned long pixel; unsigned short red, green, blue; char flags, pad; } XColor;
typedef union _XEventStub {
  int type;
  struct { int type; union { long l[5]; } data; } xclient;
} XEvent;
enum { KeyPress = 2, Expose = 12, ClientMessage = 33 };
enum { False = 0, True = 1 };
static const long KeyPressMask = 1L << 0;
static const long ExposureMask = 1L << 15;
static inline int XInitThreads(void) { return 1; }
static inline Display *XOpenDisplay(const char *) { return 0; }
static inline int DefaultScreen(Display *) { return 0; }
static inline Window RootWindow(Display *, int) { return 0; }
static inline unsigned long BlackPixel(Display *, int) { return 0; }
static inline unsigned long WhitePixel(Display *, int) { return 0; }
static inline Colormap DefaultColormap(Display *, int) { return 0; }
static inline Window XCreateSimpleWindow(Display *, Window, int, int, unsigned, unsigned, unsigned,
                                         unsigned long, unsigned long) { return 0; }
static inline int XSelectInput(Display *, Window, long) { return 0; }
static inline int XMapWindow(Display *, Window) { return 0; }
static inline GC XCreateGC(Display *, Window, unsigned long, void *) { return 0; }
static inline Atom XInternAtom(Display *, const char *, int) { return 0; }
static inline int XSetWMProtocols(Display *, Window, Atom *, int) { return 0; }
static inline int XAllocColor(Display *, Colormap, XColor *) { return 0; }
static inline int XSetForeground(Display *, GC, unsigned long) { return 0; }
static inline int XDrawPoint(Display *, Window, GC, int, int) { return 0; }
static inline int XFlush(Display *) { return 0; }
static inline int XNextEvent(Display *, XEvent *) { return 0; }
static inline int XFreeGC(Display *, GC) { return 0; }
static inline int XDestroyWindow(Display *, Window) { return 0; }
static inline int XCloseDisplay(Display *) { return 0; }
#endif
