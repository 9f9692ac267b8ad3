"""ctypes binding of libdrynx_gpu.so (include/drynx_gpu.h). Fails loudly when the library or a B200 is missing:
there is no CPU fallback anywhere in this package."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('DRYNX_LIB', os.path.join(_HERE, 'libdrynx_gpu.so'))  # DRYNX_LIB: experiment builds only

c_u8p = ctypes.POINTER(ctypes.c_uint8)
c_i64p = ctypes.POINTER(ctypes.c_int64)
c_i32p = ctypes.POINTER(ctypes.c_int32)
c_u32p = ctypes.POINTER(ctypes.c_uint32)
c_szp = ctypes.POINTER(ctypes.c_size_t)
c_dp = ctypes.POINTER(ctypes.c_double)
VP = ctypes.c_void_p

# name -> (restype, argtypes); every symbol include/drynx_gpu.h declares
SIGNATURES = {
    'drx_init': (ctypes.c_int, [ctypes.POINTER(VP), ctypes.c_int, VP, VP]),
    'drx_free': (None, [VP]),
    'drx_last_error': (ctypes.c_char_p, []),
    'drx_set_stream': (ctypes.c_int, [VP, VP]),
    'drx_synchronize': (ctypes.c_int, [VP]),
    'drx_take_bad': (ctypes.c_int, [VP, ctypes.POINTER(ctypes.c_int)]),
    'drx_launch_count': (ctypes.c_uint64, [VP]),
    'drx_set_option': (ctypes.c_int, [VP, ctypes.c_char_p, ctypes.c_int]),
    'drx_profile_enable': (ctypes.c_int, [VP, ctypes.c_int]),
    'drx_profile_get': (ctypes.c_int, [VP, ctypes.c_char_p, c_dp, ctypes.POINTER(ctypes.c_uint64)]),
    'drx_reduce_rows_dev': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_size_t, VP, ctypes.c_int, VP]),
    'drx_ed_set_pubkey': (ctypes.c_int, [VP, VP]),
    'drx_ed_encrypt': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_ed_encrypt_dev': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_ed_aggregate': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_size_t, VP, VP]),
    'drx_ed_aggregate_dev': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_size_t, VP, VP]),
    'drx_ed_dlog_table': (ctypes.c_int, [VP, ctypes.c_int64]),
    'drx_ed_decrypt': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, ctypes.c_int, VP, VP]),
    'drx_ed_decrypt_dev': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_ed_point_mul': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_freq_count': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_size_t, VP, ctypes.c_int64, ctypes.c_int64, VP]),
    'drx_logreg_coefficients': (ctypes.c_int, [VP, ctypes.c_size_t, VP, ctypes.c_int, ctypes.c_int, VP, VP, VP, VP, ctypes.c_double, VP]),
    'drx_logreg_coefficients_dev': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_size_t, VP, ctypes.c_int, ctypes.c_int, VP, VP, VP, VP, ctypes.c_double, VP]),
    'drx_encrypt_scalar': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_reduce_rows': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_size_t, VP, ctypes.c_int, VP]),
    'drx_set_pubkey': (ctypes.c_int, [VP, VP]),
    'drx_encrypt': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_encrypt_dev': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_aggregate': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_size_t, VP, VP]),
    'drx_aggregate_dev': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_size_t, VP, VP]),
    'drx_aggregate_noalloc_dev': (ctypes.c_int, [ctypes.c_size_t, ctypes.c_size_t, VP, VP, VP]),
    'drx_dlog_table': (ctypes.c_int, [VP, ctypes.c_int64]),
    'drx_decrypt': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, ctypes.c_int, VP, VP]),
    'drx_decrypt_dev': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_ct_scalar_mul': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_ct_scalar_mul_dev': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_shuffle_rerandomize': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_size_t, VP, VP, VP, VP]),
    'drx_shuffle_rerandomize_dev': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_size_t, VP, VP, VP, VP]),
    'drx_schnorr_verify': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP, VP]),
    'drx_rep_commit': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_int, VP, VP, VP]),
    'drx_rep_commit_dev': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_int, VP, VP, VP]),
    'drx_rep_respond': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_int, VP, VP, VP, VP]),
    'drx_rep_verify': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_int, VP, VP, VP, VP, VP, VP]),
    'drx_rep_verify_dev': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_int, VP, VP, VP, VP, VP, VP]),
    'drx_set_target_pubkey': (ctypes.c_int, [VP, VP]),
    'drx_keyswitch': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP, VP, VP]),
    'drx_keyswitch_dev': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP, VP, VP]),
    'drx_decrypt_check_zero': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_rp_set_sigs': (ctypes.c_int, [VP, ctypes.c_int, ctypes.c_int, ctypes.c_int64, VP, VP]),
    'drx_rp_make_sigs': (ctypes.c_int, [VP, ctypes.c_int, ctypes.c_int64, VP, VP, VP]),
    'drx_rp_ndigits': (ctypes.c_int64, [ctypes.c_int64, ctypes.c_int64, ctypes.c_int64]),
    'drx_rp_proof_size': (ctypes.c_size_t, [ctypes.c_int, ctypes.c_int64]),
    'drx_rp_create': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP, VP, VP, VP, VP, VP, VP]),
    'drx_rp_verify': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP, VP, VP]),
    'drx_rp_verify_dev': (ctypes.c_int, [VP, ctypes.c_size_t, VP, ctypes.c_int, ctypes.c_int64, VP, VP]),
    'drx_rp_create_dev': (ctypes.c_int, [VP, ctypes.c_size_t, VP, ctypes.c_int, ctypes.c_int64, VP, VP, VP, VP, VP]),
    'drx_pair': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_g1_mul': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_g2_mul': (ctypes.c_int, [VP, ctypes.c_size_t, VP, VP, VP]),
    'drx_mont_mul': (ctypes.c_int, [VP, ctypes.c_size_t, ctypes.c_int, VP, VP, VP]),
    'drx_microbench': (ctypes.c_int, [VP, ctypes.c_int, ctypes.c_int, c_dp, c_dp]),
}

_lib = None


class DrynxGpuError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__('libdrynx_gpu error %d: %s' % (code, msg))
        self.code = code


def load():
    """Loads the shared library (symbols only; no GPU needed until drx_init)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError('%s is missing: run `python -m drynx_b200.build gpu` (there is no CPU fallback)' % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the ABI lost a symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise DrynxGpuError(rc, load().drx_last_error().decode('utf-8', 'replace'))
