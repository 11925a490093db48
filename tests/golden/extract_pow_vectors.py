"""Extracts the reference's own proof-of-work vectors into tests/golden/pow_vectors.json (run in the build container,
where /root/reference exists; the JSON is committed, the reference sources are not copied):
  * src/test/arith_uint256_tests.cpp  bignum_SetCompact: (compact -> hex, GetCompact, fNegative, fOverflow)
  * src/test/pow_tests.cpp            get_next_work*: CalculateNextWorkRequired cases (height, times, bits -> bits)
                                      cash_plus_difficulty_test: the checkpoints of the scripted sequence
"""
import json
import os
import re
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def compact_vectors():
    src = open(os.path.join(REF, "src/test/arith_uint256_tests.cpp")).read()
    body = src[src.index("BOOST_AUTO_TEST_CASE(bignum_SetCompact)"):]
    body = body[:body.index("BOOST_AUTO_TEST_CASE", 10)] if "BOOST_AUTO_TEST_CASE" in body[10:] else body
    out = []
    chunks = re.split(r"num\.SetCompact\(", body)[1:]
    for ch in chunks:
        c = int(re.match(r"(0x[0-9a-fA-F]+|0)", ch).group(1), 16)
        v = {"compact": c}
        m = re.search(r'"([0-9a-f]{64})"', ch)
        if m:
            v["hex"] = m.group(1)
        m = re.search(r"GetCompact\((true)?\),\s*(0x[0-9a-fA-F]+|0)U", ch)
        if m:
            v["get_compact"] = int(m.group(2), 16)
            v["get_compact_negative_arg"] = bool(m.group(1))
        m = re.search(r"fNegative,\s*(true|false)", ch)
        if m:
            v["negative"] = m.group(1) == "true"
        m = re.search(r"fOverflow,\s*(true|false)", ch)
        if m:
            v["overflow"] = m.group(1) == "true"
        out.append(v)
    return out


def legacy_vectors():
    src = open(os.path.join(REF, "src/test/pow_tests.cpp")).read()
    out = []
    for m in re.finditer(r"nLastRetargetTime = (\d+);.*?nHeight = (\d+);.*?nTime = (\d+);.*?nBits = (0x[0-9a-f]+);.*?config\),\s*(0x[0-9a-f]+)\);",
                         src, re.S):
        out.append({"first_time": int(m.group(1)), "height": int(m.group(2)), "time": int(m.group(3)),
                    "bits": int(m.group(4), 16), "expect": int(m.group(5), 16)})
    return out


def cashplus_checkpoints():
    src = open(os.path.join(REF, "src/test/pow_tests.cpp")).read()
    body = src[src.index("BOOST_AUTO_TEST_CASE(cash_plus_difficulty_test)"):]
    return [int(x, 16) for x in re.findall(r"BOOST_CHECK_EQUAL\(nBits, (0x[0-9a-f]+)\);", body)]


if __name__ == "__main__":
    d = {"source": "src/test/arith_uint256_tests.cpp (bignum_SetCompact), src/test/pow_tests.cpp",
         "set_compact": compact_vectors(), "calculate_next_work": legacy_vectors(),
         "cashplus_checkpoints": cashplus_checkpoints()}
    assert len(d["set_compact"]) >= 20 and len(d["calculate_next_work"]) == 4 and len(d["cashplus_checkpoints"]) == 6, \
        (len(d["set_compact"]), len(d["calculate_next_work"]), d["cashplus_checkpoints"])
    json.dump(d, open(os.path.join(HERE, "pow_vectors.json"), "w"), indent=1)
    print(len(d["set_compact"]), "compact vectors,", len(d["calculate_next_work"]), "legacy retarget vectors,",
          [hex(x) for x in d["cashplus_checkpoints"]])
