"""oracle/pow_oracle.py — TEST INFRASTRUCTURE ONLY.  Python-integer restatement of the reference's proof-of-work
arithmetic and difficulty rules (row N4), used by tests/ as the checker for the product's C++ (bcp_headers.h) and
CUDA (pow.cuh) code.  Pinned by the reference's own vectors (src/test/pow_tests.cpp, src/test/arith_uint256_tests.cpp
bignum_SetCompact) and by the reference's arith_uint256 compiled in place (oracle/_ref, ref_* functions).

  set_compact / get_compact   arith_uint256::SetCompact / GetCompact   src/arith_uint256.cpp:181-232
  block_proof                 GetBlockProof                            src/chain.cpp:132-145
  check_pow                   CheckProofOfWork                         src/pow.cpp:141-161
  Chain.next_work             GetNextWorkRequired                      src/pow.cpp:72-105
  Chain.legacy_next_work      BitcoinGetNextWorkRequired               src/pow.cpp:27-67
  Chain.calculate_next_work   CalculateNextWorkRequired                src/pow.cpp:110-134
  Chain.cashplus_next_work    GetNextCashPlusWorkRequired              src/pow.cpp:252-295
  Chain.median_time_past      CBlockIndex::GetMedianTimePast           src/chain.h:312-325
"""
M256 = (1 << 256) - 1


def set_compact(c):
    size, word = c >> 24, c & 0x007FFFFF
    if size <= 3:
        word >>= 8 * (3 - size)
        t = word
    else:
        t = (word << (8 * (size - 3))) & M256  # base_uint <<= drops what leaves 256 bits
    neg = word != 0 and (c & 0x00800000) != 0
    ovf = word != 0 and (size > 34 or (word > 0xFF and size > 33) or (word > 0xFFFF and size > 32))
    return t, neg, ovf


def get_compact(t):
    size = (t.bit_length() + 7) // 8
    if size <= 3:
        c = (t & 0xFFFFFFFFFFFFFFFF) << (8 * (3 - size))
    else:
        c = (t >> (8 * (size - 3))) & 0xFFFFFFFFFFFFFFFF
    c &= 0xFFFFFFFF  # assignment to uint32_t
    if c & 0x00800000:
        c >>= 8
        size += 1
    return c | (size << 24)


def block_proof(nbits):
    t, neg, ovf = set_compact(nbits)
    if neg or ovf or t == 0:
        return 0
    return ((~t & M256) // (t + 1)) + 1


def check_pow(hash_le32, nbits, limit):
    t, neg, ovf = set_compact(nbits)
    if neg or t == 0 or ovf or t > limit:
        return False
    return int.from_bytes(hash_le32, "little") <= t


class Params:
    def __init__(self, pow_limit, pow_limit_legacy, pow_limit_start, allow_min_difficulty=False, no_retargeting=False,
                 target_spacing=600, target_timespan_legacy=14 * 24 * 3600, bcp_height=0, premine_window=0, averaging_window=30):
        self.__dict__.update(locals())

    def limit(self, postfork):
        return self.pow_limit if postfork else self.pow_limit_legacy

    @classmethod
    def main(cls):
        return cls((1 << 255) - 1, (1 << 224) - 1, (1 << 236) - 1, bcp_height=509696, premine_window=16000)


class Chain:
    """times/bits by height; chain work accumulated like CBlockIndex::nChainWork"""

    def __init__(self, params):
        self.p, self.times, self.bits, self.work = params, [], [], []

    def add(self, time, bits):
        self.times.append(time)
        self.bits.append(bits)
        self.work.append(((self.work[-1] if self.work else 0) + block_proof(bits)) & M256)

    def median_time_past(self, h=None):
        h = len(self.times) - 1 if h is None else h
        t = sorted(self.times[max(0, h - 10):h + 1])
        return t[len(t) // 2]

    def calculate_next_work(self, first_time):
        p, prev = self.p, len(self.times) - 1
        if p.no_retargeting:
            return self.bits[prev]
        span = self.times[prev] - first_time
        span = max(span, p.target_timespan_legacy // 4)
        span = min(span, p.target_timespan_legacy * 4)
        n = (set_compact(self.bits[prev])[0] * (span & 0xFFFFFFFF)) & M256  # operator*=(uint32_t)
        n //= p.target_timespan_legacy
        return get_compact(min(n, p.pow_limit_legacy))

    def legacy_next_work(self, new_time):
        p, prev = self.p, len(self.times) - 1
        limit = get_compact(p.pow_limit_legacy)
        interval = p.target_timespan_legacy // p.target_spacing
        if (prev + 1) % interval != 0:
            if p.allow_min_difficulty:
                if new_time > self.times[prev] + 2 * p.target_spacing:
                    return limit
                i = prev
                while i > 0 and i % interval != 0 and self.bits[i] == limit:
                    i -= 1
                return self.bits[i]
            return self.bits[prev]
        return self.calculate_next_work(self.times[prev - (interval - 1)])

    def _suitable(self, h):
        b = [h - 2, h - 1, h]
        t = self.times
        if t[b[0]] > t[b[2]]:
            b[0], b[2] = b[2], b[0]
        if t[b[0]] > t[b[1]]:
            b[0], b[1] = b[1], b[0]
        if t[b[1]] > t[b[2]]:
            b[1], b[2] = b[2], b[1]
        return b[1]

    def cashplus_next_work(self, new_time=0):
        p, prev = self.p, len(self.times) - 1
        postfork = prev >= p.bcp_height
        if p.allow_min_difficulty and new_time > self.times[prev] + 2 * p.target_spacing:
            return get_compact(p.limit(postfork))
        last, first = self._suitable(prev), self._suitable(prev - 144)
        work = ((self.work[last] - self.work[first]) & M256) * (p.target_spacing & 0xFFFFFFFF) & M256
        span = self.times[last] - self.times[first]
        span = min(span, 288 * p.target_spacing)
        span = max(span, 72 * p.target_spacing)
        work //= span
        nxt = ((-work) & M256) // work
        return get_compact(min(nxt, p.limit(postfork)))

    def next_work(self, new_time):
        p = self.p
        if not self.times:
            return get_compact(p.pow_limit_legacy)
        prev = len(self.times) - 1
        if p.no_retargeting:
            return self.bits[prev]
        h = prev + 1
        if h < p.bcp_height:
            return self.legacy_next_work(new_time)
        if h < p.bcp_height + p.premine_window:
            return get_compact(p.pow_limit)
        if h < p.bcp_height + p.premine_window + p.averaging_window:
            return get_compact(p.pow_limit_start)
        return self.cashplus_next_work(new_time)
