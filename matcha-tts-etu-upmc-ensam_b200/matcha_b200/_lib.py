"""ctypes binding of libmatcha_b200.so (the C ABI in include/matcha_b200.h)."""
import ctypes
import os

_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_LIB = None


class LibraryNotBuilt(RuntimeError):
    pass


class MatchaB200Error(RuntimeError):
    def __init__(self, code, where):
        lib = load_library()
        msg = lib.matcha_b200_error_string(code).decode()
        cuda = lib.matcha_b200_last_cuda_error()
        super().__init__(f"{where}: {msg} (code {code}, last CUDA error {cuda})")
        self.code = code


def lib_path():
    # MB200_LIB: an alternative build of the same library (A/B timing of a compile-time switch); never a fallback
    return os.environ.get("MB200_LIB") or os.path.join(_PKG_ROOT, "lib", "libmatcha_b200.so")


def load_library():
    """Loads the CUDA library; raises LibraryNotBuilt if it has not been compiled."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise LibraryNotBuilt(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  matcha_b200 has no CPU fallback.")
    lib = ctypes.CDLL(path)
    c_int, c_void_p, c_float, c_char_p = ctypes.c_int, ctypes.c_void_p, ctypes.c_float, ctypes.c_char_p
    lib.matcha_b200_error_string.restype = c_char_p
    lib.matcha_b200_error_string.argtypes = [c_int]
    lib.matcha_b200_last_cuda_error.restype = c_int
    lib.matcha_b200_version.restype = c_char_p
    lib.matcha_b200_mas_compute_batch_alignments.restype = c_int
    lib.matcha_b200_mas_compute_batch_alignments.argtypes = [
        c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
        c_int, c_int, c_int, c_float, c_void_p, c_void_p]
    lib.matcha_b200_mas_lengths_from_mask.restype = c_int
    lib.matcha_b200_mas_lengths_from_mask.argtypes = [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]
    lib.matcha_b200_mas_maximum_path_host.restype = c_int
    lib.matcha_b200_mas_maximum_path_host.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float]
    lib.matcha_b200_mas_host_create.restype = c_void_p
    lib.matcha_b200_mas_host_destroy.restype = c_int
    lib.matcha_b200_mas_host_destroy.argtypes = [c_void_p]
    lib.matcha_b200_mas_maximum_path_host_h.restype = c_int
    lib.matcha_b200_mas_maximum_path_host_h.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float]
    lib.matcha_b200_mas_host_release.restype = c_int
    lib.matcha_b200_mas_query_launch.restype = c_int
    lib.matcha_b200_mas_query_launch.argtypes = [c_int, c_int, c_int, c_int, c_int, ctypes.POINTER(c_int * 8)]
    lib.matcha_b200_mas_query_path.restype = c_int
    lib.matcha_b200_mas_query_path.argtypes = [c_int, c_int, c_int, c_int, c_int]
    lib.matcha_b200_mas_apply_mask.restype = c_int
    lib.matcha_b200_mas_apply_mask.argtypes = [c_void_p, c_void_p, c_void_p, ctypes.c_longlong, c_void_p]
    ll = ctypes.c_longlong
    lib.matcha_b200_mas_align.restype = c_int
    lib.matcha_b200_mas_align.argtypes = [
        c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
        c_int, c_int, c_int, c_float, c_void_p, c_void_p]
    lib.matcha_b200_mas_durations_static_max_cells.restype = ll
    lib.matcha_b200_mas_durations_static.restype = c_int
    lib.matcha_b200_mas_durations_static.argtypes = [
        c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
        c_int, c_int, c_int, c_float, c_void_p, c_void_p]
    lib.matcha_b200_align_log_prior.restype = c_int
    lib.matcha_b200_align_log_prior.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                                c_int, c_int, c_int, c_int, c_void_p]
    lib.matcha_b200_align_expand.restype = c_int
    lib.matcha_b200_align_expand.argtypes = [c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                             c_void_p, c_void_p, c_void_p, c_void_p,
                                             c_int, c_int, c_int, c_int, c_void_p]
    lib.matcha_b200_cfm_loss_mix.restype = c_int
    lib.matcha_b200_cfm_loss_mix.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, ctypes.c_double, c_int, ll, c_void_p]
    lib.matcha_b200_cfm_loss_mse.restype = c_int
    lib.matcha_b200_cfm_loss_mse.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, ctypes.c_double,
                                             ll, ll, c_void_p]
    lib.matcha_b200_decoder_blob_floats.restype = ll
    lib.matcha_b200_decoder_workspace_bytes.restype = ll
    lib.matcha_b200_decoder_workspace_bytes.argtypes = [c_int, c_int, c_int, c_int]
    lib.matcha_b200_decoder_blob_bf16_bytes.restype = ll
    lib.matcha_b200_decoder_pack_bf16.restype = c_int
    lib.matcha_b200_decoder_pack_bf16.argtypes = [c_void_p, c_void_p, c_int, c_void_p]
    lib.matcha_b200_decoder_forward.restype = c_int
    lib.matcha_b200_decoder_forward.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                                                c_void_p, ll, c_int, c_int, c_int, c_void_p]
    lib.matcha_b200_cfm_loss_forward.restype = c_int
    lib.matcha_b200_cfm_loss_forward.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, ctypes.c_double, c_void_p, c_void_p,
                                                 c_void_p, c_void_p, c_void_p, c_void_p, ll, c_int, c_int, c_int, c_void_p]
    lib.matcha_b200_cfm_solve_euler.restype = c_int
    lib.matcha_b200_cfm_solve_euler.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                                c_void_p, c_void_p, c_void_p, ll, c_int, c_int, c_int, c_void_p]
    lib.matcha_b200_text_encoder_blob_floats.restype = ll
    lib.matcha_b200_text_encoder_blob_floats.argtypes = [c_int]
    lib.matcha_b200_text_encoder_workspace_bytes.restype = ll
    lib.matcha_b200_text_encoder_workspace_bytes.argtypes = [c_int, c_int, c_int]
    lib.matcha_b200_text_encoder_forward.restype = c_int
    lib.matcha_b200_text_encoder_forward.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                                     c_void_p, c_int, c_int, c_int, c_void_p, ll, c_void_p]
    lib.matcha_b200_debug_attention.restype = c_int
    lib.matcha_b200_debug_attention.argtypes = [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]
    lib.matcha_b200_debug_gemm_bf16.restype = c_int
    lib.matcha_b200_debug_gemm_bf16.argtypes = [c_void_p, c_int, c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_void_p,
                                                c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]
    _LIB = lib
    return lib


def check(code, where):
    if code < 0:
        raise MatchaB200Error(code, where)
    return code


# every symbol include/matcha_b200.h declares (tests check the .so exports all of them)
DECLARED_SYMBOLS = (
    "matcha_b200_error_string",
    "matcha_b200_last_cuda_error",
    "matcha_b200_version",
    "matcha_b200_mas_compute_batch_alignments",
    "matcha_b200_mas_lengths_from_mask",
    "matcha_b200_mas_maximum_path_host",
    "matcha_b200_mas_host_create",
    "matcha_b200_mas_host_destroy",
    "matcha_b200_mas_maximum_path_host_h",
    "matcha_b200_mas_host_release",
    "matcha_b200_mas_query_launch",
    "matcha_b200_mas_query_path",
    "matcha_b200_mas_apply_mask",
    "matcha_b200_mas_align",
    "matcha_b200_mas_durations_static_max_cells",
    "matcha_b200_mas_durations_static",
    "matcha_b200_align_log_prior",
    "matcha_b200_align_expand",
    "matcha_b200_cfm_loss_mix",
    "matcha_b200_cfm_loss_mse",
    "matcha_b200_decoder_blob_floats",
    "matcha_b200_decoder_workspace_bytes",
    "matcha_b200_decoder_blob_bf16_bytes",
    "matcha_b200_decoder_pack_bf16",
    "matcha_b200_decoder_forward",
    "matcha_b200_cfm_loss_forward",
    "matcha_b200_cfm_solve_euler",
    "matcha_b200_text_encoder_blob_floats",
    "matcha_b200_text_encoder_workspace_bytes",
    "matcha_b200_text_encoder_forward",
)
