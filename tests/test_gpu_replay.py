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
