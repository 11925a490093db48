"""ctypes face over bcp::HeaderChain (include/compat/bcp_headers.h, built into libbcp_chain.so): the host side of
the header pipeline — difficulty, median time past, headers-first acceptance."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))


class PowParams(C.Structure):
    """Plain-C mirror of bcp::PowParams (the Consensus::Params fields the header rules read,
    src/consensus/params.h:40-93); 256-bit values as uint256 memory bytes (little-endian)."""
    _fields_ = [("pow_limit", C.c_uint8 * 32), ("pow_limit_legacy", C.c_uint8 * 32), ("pow_limit_start", C.c_uint8 * 32),
                ("allow_min_difficulty", C.c_int32), ("no_retargeting", C.c_int32),
                ("target_spacing", C.c_int64), ("target_timespan_legacy", C.c_int64),
                ("bcp_height", C.c_int32), ("premine_window", C.c_int32), ("averaging_window", C.c_int64),
                ("bip34_height", C.c_int32), ("bip65_height", C.c_int32), ("bip66_height", C.c_int32),
                ("eh_n", C.c_uint32), ("eh_k", C.c_uint32)]

    @classmethod
    def make(cls, pow_limit, pow_limit_legacy, pow_limit_start, **kw):
        p = cls()
        for name, v in (("pow_limit", pow_limit), ("pow_limit_legacy", pow_limit_legacy), ("pow_limit_start", pow_limit_start)):
            C.memmove(getattr(p, name), int(v).to_bytes(32, "little"), 32)
        p.target_spacing, p.target_timespan_legacy, p.averaging_window = 600, 14 * 24 * 3600, 30
        p.eh_n, p.eh_k = 200, 9
        for k, v in kw.items():
            setattr(p, k, v)
        return p

    @classmethod
    def main(cls):  # src/chainparams.cpp:98-160
        return cls.make((1 << 255) - 1, (1 << 224) - 1, (1 << 236) - 1, bcp_height=509696, premine_window=16000,
                        bip34_height=227931, bip65_height=388381, bip66_height=363725)

    @classmethod
    def regtest(cls):  # src/chainparams.cpp:335-392
        lim = (1 << 255) - 1
        return cls.make(lim, lim, lim, bcp_height=3000, premine_window=20, allow_min_difficulty=1, no_retargeting=1,
                        bip34_height=100000000, bip65_height=1351, bip66_height=1251, eh_n=48, eh_k=5)


_L = None


def chain_lib():
    global _L
    if _L is None:
        L = C.CDLL(os.path.join(_HERE, "libbcp_chain.so"))
        vp, u32, i64 = C.c_void_p, C.c_uint32, C.c_int64
        sigs = {
            "bcp_headers_new": (vp, [C.POINTER(PowParams)]),
            "bcp_headers_free": (None, [vp]),
            "bcp_headers_add_index": (None, [vp, C.c_int32, u32, u32, C.c_char_p]),
            "bcp_headers_height": (C.c_int, [vp]),
            "bcp_headers_next_work": (u32, [vp, i64]),
            "bcp_headers_calculate_next_work": (u32, [vp, i64]),
            "bcp_headers_cashplus_next_work": (u32, [vp, i64]),
            "bcp_headers_median_time_past": (i64, [vp]),
            "bcp_headers_chain_work": (None, [vp, C.c_char_p]),
            "bcp_headers_accept": (C.c_long, [vp, C.c_int, vp, C.c_size_t, C.c_size_t, C.c_size_t, i64, C.c_char_p, C.c_size_t,
                                              C.POINTER(C.c_long)]),
            "bcp_compact_to_target": (C.c_int, [u32, C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
            "bcp_compact_from_target": (u32, [C.c_char_p]),
            "bcp_block_proof": (None, [u32, C.c_char_p]),
        }
        for name, (res, args) in sigs.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _L = L
    return _L


def compact_to_target(compact):
    out, neg, ovf = C.create_string_buffer(32), C.c_int(0), C.c_int(0)
    chain_lib().bcp_compact_to_target(compact, out, C.byref(neg), C.byref(ovf))
    return int.from_bytes(out.raw, "little"), bool(neg.value), bool(ovf.value)


def compact_from_target(target):
    return chain_lib().bcp_compact_from_target(int(target).to_bytes(32, "little"))


def block_proof(nbits):
    out = C.create_string_buffer(32)
    chain_lib().bcp_block_proof(nbits, out)
    return int.from_bytes(out.raw, "little")


class HeaderChain:
    def __init__(self, params):
        self.params = params
        self.h = chain_lib().bcp_headers_new(C.byref(params))

    def __del__(self):
        if getattr(self, "h", None):
            chain_lib().bcp_headers_free(self.h)
            self.h = None

    def add_index(self, version, time, bits, hash32=b"\0" * 32):
        chain_lib().bcp_headers_add_index(self.h, version, time, bits, bytes(hash32))

    @property
    def height(self):
        return chain_lib().bcp_headers_height(self.h)

    def next_work(self, new_time):
        return chain_lib().bcp_headers_next_work(self.h, new_time)

    def calculate_next_work(self, first_block_time):
        return chain_lib().bcp_headers_calculate_next_work(self.h, first_block_time)

    def cashplus_next_work(self, new_time=0):
        return chain_lib().bcp_headers_cashplus_next_work(self.h, new_time)

    def median_time_past(self):
        return chain_lib().bcp_headers_median_time_past(self.h)

    def chain_work(self):
        out = C.create_string_buffer(32)
        chain_lib().bcp_headers_chain_work(self.h, out)
        return int.from_bytes(out.raw, "little")

    def accept(self, device, hdrs, adjusted_time):
        """hdrs: (count, hdr_len) uint8 array -> (accepted, reason, failed_index)"""
        import numpy as np

        hdrs = np.ascontiguousarray(hdrs, dtype=np.uint8)
        count, hdr_len = hdrs.shape
        reason, failed = C.create_string_buffer(256), C.c_long(-1)
        n = chain_lib().bcp_headers_accept(self.h, device, hdrs.ctypes.data, hdr_len, hdr_len, count, adjusted_time, reason, 256,
                                           C.byref(failed))
        return n, reason.value.decode(), failed.value
