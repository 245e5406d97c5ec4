"""ctypes mirror of include/b200jpeg.h (the C ABI of the scan-to-pixels path).

Plumbing only: the structs here are byte-for-byte the ones declared in the header, which in turn
mirror the reference's HuffmanTable (huffman.hpp:8-15), QuantizationTable (types.hpp:79-96),
ColorComponent (types.hpp:69-77) and Image's private state (jpeg.cpp:21-34).
"""
import ctypes as C

ABI_VERSION = 2

# status bits (b200jpeg.h)
ST_BAD_MARKER = 1 << 0
ST_NO_EOI = 1 << 1
ST_BAD_RST_COUNT = 1 << 2
ST_BAD_CODE = 1 << 3
ST_BAD_DC_LEN = 1 << 4
ST_BAD_RUN = 1 << 5
ST_BAD_AC_LEN = 1 << 6
ST_OVERRUN = 1 << 7
ST_BLOCK_COUNT = 1 << 8
ST_DC_RANGE = 1 << 9
ST_NOT_CONVERGED = 1 << 10
ST_UNSUPPORTED = 1 << 11

E_ARG, E_CUDA, E_NOMEM, E_STATE = -1, -2, -3, -4
E_NOT_JPEG, E_PROGRESSIVE, E_COMPONENTS, E_SAMPLING, E_TABLE_ID, E_DHT = -10, -11, -12, -13, -14, -15
E_SOS, E_DRI, E_NO_SOS, E_TABLE_MISSING, E_TRUNCATED, E_COMPONENT_ID, E_UNSUPPORTED = -16, -17, -18, -19, -20, -21, -22
E_TOO_LARGE = -23
NOALLOC_MAX_BYTES = 16 << 20

PARSE_REFERENCE, PARSE_SPEC = 0, 1
ENTROPY_AUTO, ENTROPY_SEGMENT, ENTROPY_SYNC = 0, 1, 2

T_UNSTUFF, T_SYNC, T_WRITE, T_PIXEL, T_TOTAL, T_COUNT = 0, 1, 2, 3, 4, 5


class HuffmanTable(C.Structure):
    _fields_ = [("offset", C.c_uint8 * 17), ("symbols", C.c_uint8 * 176), ("codes", C.c_uint32 * 176),
                ("set", C.c_uint8)]


class QuantTable(C.Structure):
    _fields_ = [("table", C.c_int32 * 64), ("set", C.c_uint8)]


class Component(C.Structure):
    _fields_ = [("horizontal_sampling_factor", C.c_int32), ("vertical_sampling_factor", C.c_int32),
                ("quantizationTableID", C.c_int32), ("huffmanDCTTableID", C.c_int32),
                ("huffmanACTableID", C.c_int32)]


class ImageDesc(C.Structure):
    _fields_ = [("image_width", C.c_int32), ("image_height", C.c_int32),
                ("mcuWidth", C.c_int32), ("mcuHeight", C.c_int32),
                ("mcuWidthReal", C.c_int32), ("mcuHeightReal", C.c_int32),
                ("horizontalSamplingFactor", C.c_int32), ("verticalSamplingFactor", C.c_int32),
                ("numComponents", C.c_int32), ("restartInterval", C.c_int32), ("zeroBased", C.c_int32),
                ("color_components", Component * 3),
                ("quantization_tables", QuantTable * 4),
                ("huffmanDCTTable", HuffmanTable * 4), ("huffmanACTable", HuffmanTable * 4),
                ("scan_offset", C.c_int64), ("scan_length", C.c_int64)]


PROG_MAX_SCANS, PROG_MAX_TABLES = 48, 48


class Scan(C.Structure):
    _fields_ = [("ns", C.c_int32), ("comp", C.c_int32 * 3), ("dc_table", C.c_int32 * 3), ("ac_table", C.c_int32 * 3),
                ("ss", C.c_int32), ("se", C.c_int32), ("ah", C.c_int32), ("al", C.c_int32), ("restart_interval", C.c_int32),
                ("offset", C.c_int64), ("length", C.c_int64)]


class ProgressiveDesc(C.Structure):
    _fields_ = [("frame", ImageDesc), ("nscans", C.c_int32), ("ntables", C.c_int32), ("scans", Scan * PROG_MAX_SCANS),
                ("tables", HuffmanTable * PROG_MAX_TABLES)]


class BatchStats(C.Structure):
    _fields_ = [("images", C.c_int64), ("pixels", C.c_int64), ("scan_bytes", C.c_int64),
                ("coded_blocks", C.c_int64), ("rgb_bytes", C.c_int64), ("h2d_bytes", C.c_int64),
                ("d2h_bytes", C.c_int64), ("kernel_launches", C.c_int32), ("sync_rounds", C.c_int32),
                ("entropy_path", C.c_int32), ("reserved", C.c_int32)]


# every symbol include/b200jpeg.h declares
EXPORTS = [
    "b200jpeg_parse_header", "b200jpeg_create", "b200jpeg_destroy", "b200jpeg_last_error",
    "b200jpeg_decode_batch", "b200jpeg_batch_prepare", "b200jpeg_batch_run", "b200jpeg_batch_fetch",
    "b200jpeg_sync", "b200jpeg_device_rgb", "b200jpeg_stream", "b200jpeg_batch_get_stats",
    "b200jpeg_set_timing", "b200jpeg_last_timing", "b200jpeg_host_alloc", "b200jpeg_host_free",
    "b200jpeg_abi_version", "b200jpeg_set_entropy_path", "b200jpeg_parse_headers", "b200jpeg_copy_floor",
    "b200jpeg_multi_create", "b200jpeg_multi_destroy", "b200jpeg_multi_count", "b200jpeg_multi_ctx",
    "b200jpeg_multi_decode_batch", "b200jpeg_decode_batch_multi", "b200jpeg_export_coefficients_noalloc",
    "b200jpeg_parse_progressive", "b200jpeg_decode_progressive",
]


def bind_parse(lib):
    lib.b200jpeg_parse_header.restype = C.c_int
    lib.b200jpeg_parse_header.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(ImageDesc)]
    lib.b200jpeg_parse_progressive.restype = C.c_int
    lib.b200jpeg_parse_progressive.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(ProgressiveDesc)]
    lib.b200jpeg_parse_headers.restype = C.c_int
    lib.b200jpeg_parse_headers.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.c_int, C.c_int, C.c_int,
                                           C.POINTER(ImageDesc), C.POINTER(C.c_int32)]


def bind(lib):
    """Declare argument/return types of every entry point on a loaded libb200jpeg.so."""
    bind_parse(lib)
    P = C.POINTER
    lib.b200jpeg_create.restype = C.c_int
    lib.b200jpeg_create.argtypes = [C.c_int, P(C.c_void_p)]
    lib.b200jpeg_destroy.restype = None
    lib.b200jpeg_destroy.argtypes = [C.c_void_p]
    lib.b200jpeg_last_error.restype = C.c_char_p
    lib.b200jpeg_last_error.argtypes = [C.c_void_p]
    lib.b200jpeg_decode_batch.restype = C.c_int
    lib.b200jpeg_decode_batch.argtypes = [C.c_void_p, P(ImageDesc), C.c_int, P(C.c_void_p), P(C.c_size_t),
                                          P(C.c_void_p), P(C.c_void_p), P(C.c_int32)]
    lib.b200jpeg_batch_prepare.restype = C.c_int
    lib.b200jpeg_batch_prepare.argtypes = [C.c_void_p, P(ImageDesc), C.c_int, P(C.c_void_p), P(C.c_size_t)]
    lib.b200jpeg_batch_run.restype = C.c_int
    lib.b200jpeg_batch_run.argtypes = [C.c_void_p]
    lib.b200jpeg_batch_fetch.restype = C.c_int
    lib.b200jpeg_batch_fetch.argtypes = [C.c_void_p, P(C.c_void_p), P(C.c_void_p), P(C.c_int32)]
    lib.b200jpeg_sync.restype = C.c_int
    lib.b200jpeg_sync.argtypes = [C.c_void_p]
    lib.b200jpeg_device_rgb.restype = C.c_int
    lib.b200jpeg_device_rgb.argtypes = [C.c_void_p, C.c_int, P(C.c_void_p), P(C.c_size_t)]
    lib.b200jpeg_stream.restype = C.c_void_p
    lib.b200jpeg_stream.argtypes = [C.c_void_p]
    lib.b200jpeg_batch_get_stats.restype = C.c_int
    lib.b200jpeg_batch_get_stats.argtypes = [C.c_void_p, P(BatchStats)]
    lib.b200jpeg_set_entropy_path.restype = C.c_int
    lib.b200jpeg_set_entropy_path.argtypes = [C.c_void_p, C.c_int]
    lib.b200jpeg_set_timing.restype = C.c_int
    lib.b200jpeg_set_timing.argtypes = [C.c_void_p, C.c_int]
    lib.b200jpeg_last_timing.restype = C.c_int
    lib.b200jpeg_last_timing.argtypes = [C.c_void_p, P(C.c_float), C.c_int]
    lib.b200jpeg_host_alloc.restype = C.c_void_p
    lib.b200jpeg_host_alloc.argtypes = [C.c_size_t]
    lib.b200jpeg_host_free.restype = None
    lib.b200jpeg_host_free.argtypes = [C.c_void_p]
    lib.b200jpeg_abi_version.restype = C.c_int
    lib.b200jpeg_abi_version.argtypes = []
    lib.b200jpeg_export_coefficients_noalloc.restype = C.c_int
    lib.b200jpeg_export_coefficients_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
    lib.b200jpeg_decode_progressive.restype = C.c_int
    lib.b200jpeg_decode_progressive.argtypes = [C.c_void_p, P(ProgressiveDesc), C.c_int, P(C.c_void_p), P(C.c_size_t),
                                                P(C.c_void_p), P(C.c_void_p), P(C.c_int32)]
    lib.b200jpeg_copy_floor.restype = C.c_int
    lib.b200jpeg_copy_floor.argtypes = [C.c_void_p, P(ImageDesc), C.c_int, P(C.c_void_p), P(C.c_size_t), P(C.c_void_p),
                                        P(C.c_float)]
    lib.b200jpeg_multi_create.restype = C.c_int
    lib.b200jpeg_multi_create.argtypes = [P(C.c_int), C.c_int, P(C.c_void_p)]
    lib.b200jpeg_multi_destroy.restype = None
    lib.b200jpeg_multi_destroy.argtypes = [C.c_void_p]
    lib.b200jpeg_multi_count.restype = C.c_int
    lib.b200jpeg_multi_count.argtypes = [C.c_void_p]
    lib.b200jpeg_multi_ctx.restype = C.c_void_p
    lib.b200jpeg_multi_ctx.argtypes = [C.c_void_p, C.c_int]
    lib.b200jpeg_multi_decode_batch.restype = C.c_int
    lib.b200jpeg_multi_decode_batch.argtypes = [C.c_void_p, P(ImageDesc), C.c_int, P(C.c_void_p), P(C.c_size_t),
                                                P(C.c_void_p), P(C.c_int32)]
    lib.b200jpeg_decode_batch_multi.restype = C.c_int
    lib.b200jpeg_decode_batch_multi.argtypes = [P(C.c_int), C.c_int, P(ImageDesc), C.c_int, P(C.c_void_p), P(C.c_size_t),
                                                P(C.c_void_p), P(C.c_int32)]
    return lib
