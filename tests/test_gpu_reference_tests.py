"""The reference's own test suite (worker_test.py:66-189), UNCHANGED, against the drop-in: the file is
unpacked from tests/golden/worker_test.fixture into a scratch directory whose rater.py is the
one-line module of INTEGRATION.md section 1, and run by pytest in a subprocess."""
import base64
import hashlib
import os
import subprocess
import sys
import zlib

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIXTURE = os.path.join(ROOT, "tests", "golden", "worker_test.fixture")
ONE_LINE_RATER = ("# rater.py -- replaces the reference module of the same name (INTEGRATION.md section 1)\n"
                  "from analyzer_b200.rater import *          # rate_match, get_trueskill_seed, vst_points, env, TAU, ...\n"
                  "from analyzer_b200.rater import rate_matches, rate_batch, logger\n")


def unpack(dst_dir):
    lines = open(FIXTURE).read().splitlines()
    sha = lines[1].split()[2]
    raw = zlib.decompress(base64.b85decode("".join(lines[2:])))
    assert hashlib.sha256(raw).hexdigest() == sha
    with open(os.path.join(dst_dir, "worker_test.py"), "wb") as f:
        f.write(raw)
    with open(os.path.join(dst_dir, "rater.py"), "w") as f:
        f.write(ONE_LINE_RATER)
    return raw


def test_fixture_is_the_reference_file(tmp_path):
    """Here (where /root/reference exists) the fixture must be the reference's file, byte for byte."""
    raw = unpack(str(tmp_path))
    ref = "/root/reference/worker_test.py"
    if not os.path.exists(ref):
        pytest.skip("no reference tree on this box")
    assert raw == open(ref, "rb").read()


@pytest.mark.gpu
def test_reference_worker_test_passes_unchanged_on_the_gpu_path(tmp_path):
    unpack(str(tmp_path))
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    out = subprocess.run([sys.executable, "-m", "pytest", "-q", "-p", "no:cacheprovider", "worker_test.py"],
                         cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=600)
    tail = (out.stdout + out.stderr)[-3000:]
    assert out.returncode == 0, tail
    assert " passed" in out.stdout and "failed" not in out.stdout, tail
