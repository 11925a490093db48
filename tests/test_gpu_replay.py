"""the reference's own script interpreter (compiled in place, oracle/_ref) driven through this
project's CCheckQueue drop-in with a recording signature checker; multisig, result-dependent scripts and the
signature-cache hook, every verdict compared with the reference's CPU checker.  The binary is built here by
`make -C oracle ref` (it needs /root/reference to build, not to run) and travels to the GPU box."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "oracle", "_ref", "test_replay_ref")

needs_exe = pytest.mark.skipif(not os.path.exists(EXE), reason="oracle/_ref/test_replay_ref not built (reference sources absent)")


@needs_exe
def test_replay_case_set_cpu():
    """The generated case set holds both valid and invalid scripts of every family (reference CPU verdicts)."""
    out = subprocess.run([EXE, "cases"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr
    _, n, _, valid = out.stdout.split()
    assert int(n) >= 200 and 0.3 * int(n) < int(valid) < 0.9 * int(n)


@needs_exe
@pytest.mark.gpu
def test_reference_interpreter_through_checkqueue_gpu():
    out = subprocess.run([EXE], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-2000:]
    assert "replay ok" in out.stdout


def _vector_file(tmp_path):
    import gzip
    import json
    import struct

    vs = json.loads(gzip.open(os.path.join(ROOT, "tests", "golden", "script_tests.json.gz")).read())
    path = tmp_path / "script_vectors.bin"
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(vs)))
        for v in vs:
            sig, spk, tx = bytes.fromhex(v["sig"]), bytes.fromhex(v["spk"]), bytes.fromhex(v["tx"])
            f.write(struct.pack("<IqB", v["flags"], v["amount"], 1 if v["err"] == 0 else 0))
            for b in (sig, spk, tx):
                f.write(struct.pack("<I", len(b)) + b)
    return str(path), len(vs)


@needs_exe
def test_script_tests_json_fixture_matches_the_reference_cpu(tmp_path):
    """tests/golden/script_tests.json.gz (the reference's src/test/data/script_tests.json as binary scripts,
    tests/golden/make_script_vectors.py) evaluated by the reference's own VerifyScript: every expectation holds."""
    path, n = _vector_file(tmp_path)
    assert n > 1000
    out = subprocess.run([EXE, "vectors-cpu", path], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]


@needs_exe
@pytest.mark.gpu
def test_script_tests_json_through_checkqueue_gpu(tmp_path):
    """The reference's own script corpus (1,108 vectors: DER / high-S / hybrid keys / NULLFAIL / FORKID / NULLDUMMY /
    multisig / P2SH / every opcode family) through the CCheckQueue drop-in with the recording checker and the GPU: each
    vector alone gives the expected verdict; all verifying vectors form an accepted block; adding one failing vector that
    reaches a signature comparison makes the block fail."""
    path, _ = _vector_file(tmp_path)
    out = subprocess.run([EXE, "vectors", path], capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-2000:]
    assert "vectors ok" in out.stdout


MIXED = os.path.join(ROOT, "oracle", "_ref", "test_mixed_ref")


@pytest.mark.skipif(not os.path.exists(MIXED), reason="oracle/_ref/test_mixed_ref not built (reference sources absent)")
@pytest.mark.gpu
def test_mixed_blocks_connect_through_both_connectors_gpu():
    """Blocks mixing P2PKH, pay-to-pubkey, bare multisig, P2SH multisig and P2SH-P2PKH inputs: the device fast path takes the
    P2PKH inputs, everything else goes through the OtherScripts callback = the reference's interpreter inside the check
    queue drop-in (signatures on the GPU); verdicts (valid block, corrupted multisig signature, wrong signature order, bad
    P2PKH signature) equal the reference's CPU VerifyScript over all inputs; refusals leave the coin table untouched."""
    out = subprocess.run([MIXED], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-2000:]
    assert "mixed ok" in out.stdout
