"""Loader for libbcp_b200.so (in-tree build product).  Fails loudly: no fallback of any kind."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# BCP_B200_LIB lets tools/quick_bench.py A/B-test alternative builds of the same C ABI; the default is the in-tree build
_SO = os.environ.get("BCP_B200_LIB") or os.path.join(HERE, "libbcp_b200.so")


class BcpError(RuntimeError):
    pass


def lib_path():
    return _SO


_lib = None

u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
vp = C.c_void_p
sz = C.c_size_t

# name -> (restype, argtypes): exactly the declarations of include/bcp_gpu.h
SIGNATURES = {
    "bcp_gpu_device_count": (C.c_int, []),
    "bcp_gpu_peer_access": (C.c_int, [C.c_int, C.c_int]),
    "bcp_gpu_init": (C.c_int, [C.c_int]),
    "bcp_gpu_shutdown": (C.c_int, [C.c_int]),
    "bcp_gpu_last_error": (C.c_char_p, []),
    "bcp_gpu_host_alloc": (vp, [sz]),
    "bcp_gpu_host_free": (None, [vp]),
    "bcp_gpu_ecdsa_verify_batch": (C.c_int, [C.c_int, vp, vp, vp, sz, C.c_uint32, vp]),
    "bcp_gpu_ecdsa_verify_batch_dev": (C.c_int, [C.c_int, vp, vp, vp, sz, C.c_uint32, vp, vp]),
    "bcp_gpu_ecdsa_scratch_bytes": (sz, [sz]),
    "bcp_gpu_ecdsa_reserve": (C.c_int, [C.c_int, sz]),
    "bcp_gpu_pubkey_parse_batch": (C.c_int, [C.c_int, vp, sz, vp, vp]),
    "bcp_gpu_ecdsa_recover_batch": (C.c_int, [C.c_int, vp, vp, vp, sz, vp, vp]),
    "bcp_gpu_pubkey_tweak_add_batch": (C.c_int, [C.c_int, vp, vp, sz, vp, vp]),
    "bcp_der_lax_parse": (C.c_int, [vp, sz, vp]),
    "bcp_block_index_rules": (C.c_long, [vp, sz, C.c_int, vp, vp, sz, vp, vp, vp, sz, vp, vp, vp, vp, sz, vp, vp, vp, vp]),
    "bcp_block_rule_name": (C.c_char_p, [C.c_int]),
    "bcp_gpu_utxo_create": (vp, [C.c_int, C.c_uint, sz]),
    "bcp_gpu_utxo_destroy": (None, [vp]),
    "bcp_gpu_utxo_add_coins": (C.c_int, [vp, vp, vp, vp, vp, sz]),
    "bcp_gpu_utxo_size": (C.c_longlong, [vp]),
    "bcp_gpu_utxo_apply_block": (C.c_int, [vp, C.c_uint32, vp, sz, vp, sz, vp, vp, sz, vp, vp, vp, sz, vp, vp, vp, vp, sz, vp, vp]),
    "bcp_gpu_utxo_wait": (C.c_int, [vp, C.c_uint32, vp]),
    "bcp_gpu_utxo_block_scripts": (C.c_int, [vp, C.c_uint32, C.c_int, vp, vp, vp, C.c_uint32, vp, C.POINTER(C.c_int), vp]),
    "bcp_gpu_utxo_fetch_spent": (C.c_int, [vp, C.c_uint32, vp, sz, vp, vp]),
    "bcp_gpu_utxo_confirm": (C.c_int, [vp, C.c_uint32]),
    "bcp_gpu_utxo_rollback": (C.c_int, [vp, C.c_uint32]),
    "bcp_gpu_utxo_reserve": (C.c_int, [vp, C.c_size_t]),
    "bcp_block_index": (C.c_long, [vp, sz, C.c_int, vp, sz]),
    "bcp_gpu_sha256d_txids_merkle": (C.c_int, [C.c_int, vp, sz, vp, sz, vp, vp, C.POINTER(C.c_int)]),
    "bcp_gpu_merkle_root": (C.c_int, [C.c_int, vp, sz, vp, C.POINTER(C.c_int)]),
    "bcp_gpu_sha256d_batch": (C.c_int, [C.c_int, vp, sz, vp, sz, vp]),
    "bcp_gpu_sha256d_batch_dev": (C.c_int, [C.c_int, vp, vp, sz, vp, vp]),
    "bcp_gpu_merkle_scratch_bytes": (sz, [sz]),
    "bcp_gpu_merkle_root_dev": (C.c_int, [C.c_int, vp, sz, vp, vp, vp, vp]),
    "bcp_gpu_merkle_root_dev_static": (C.c_int, [C.c_int, vp, sz, vp, vp, vp]),
    "bcp_block_index_inputs": (C.c_long, [vp, sz, C.c_int, vp, sz, vp, vp, vp, sz, C.POINTER(C.c_long)]),
    "bcp_block_index_full": (C.c_long, [vp, sz, C.c_int, vp, vp, sz, vp, vp, vp, sz, C.POINTER(C.c_long), vp, vp, vp, sz,
                                        C.POINTER(C.c_long), vp]),
    "bcp_gpu_block_scripts_p2pkh": (C.c_int, [C.c_int, vp, sz, vp, sz, vp, vp, vp, sz, vp, vp, vp, C.c_uint32, vp,
                                              C.POINTER(C.c_int)]),
    "bcp_gpu_block_connect_p2pkh": (C.c_int, [C.c_int, vp, sz, C.c_int, vp, vp, vp, sz, C.c_uint32, vp, vp, C.POINTER(C.c_int),
                                              vp, C.POINTER(C.c_int)]),
    "bcp_gpu_equihash_verify_batch": (C.c_int, [C.c_int, C.c_uint, C.c_uint, vp, sz, sz, sz, vp, vp]),
    "bcp_gpu_blocks_connect_p2pkh": (C.c_int, [C.c_int, sz, vp, vp, C.c_int, vp, vp, vp, vp, C.c_uint32, vp, vp, vp, vp, vp, vp]),
    "bcp_gpu_equihash_scratch_bytes": (sz, [sz]),
    "bcp_gpu_header_check_batch": (C.c_int, [C.c_int, C.c_uint, C.c_uint, vp, sz, sz, sz, vp, vp, C.c_uint32, vp, vp, vp, vp]),
    "bcp_gpu_equihash_verify_batch_dev": (C.c_int, [C.c_int, C.c_uint, C.c_uint, vp, sz, sz, sz, vp, vp, vp]),
    "bcp_gpu_selftest": (C.c_int, [C.c_int]),
    "bcp_gpu_measure_int_peak": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_double)]),
}


def lib():
    """The loaded library with argtypes set; raises BcpError if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise BcpError(
                "%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  There is no CPU fallback." % _SO)
        try:
            L = C.CDLL(_SO)
        except OSError as e:
            raise BcpError("cannot load %s: %s" % (_SO, e))
        for name, (res, args) in SIGNATURES.items():
            try:
                f = getattr(L, name)
            except AttributeError:
                raise BcpError("%s does not export %s" % (_SO, name))
            f.restype = res
            f.argtypes = args
        _lib = L
    return _lib


def check(rc, what):
    if rc != 0:
        msg = lib().bcp_gpu_last_error() or b""
        raise BcpError("%s failed (%d): %s" % (what, rc, msg.decode("utf-8", "replace")))
