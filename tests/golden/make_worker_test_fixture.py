"""Freezes the reference's own test file as a fixture (test infrastructure only -- nothing in the
product reads it): /root/reference/worker_test.py, unmodified, zlib-compressed and base85-encoded
into tests/golden/worker_test.fixture with its SHA-256.  The reference tree does not exist on the
GPU box, so this is how its tests travel there; tests/test_gpu_reference_tests.py unpacks the file
into a scratch directory next to the one-line rater.py of INTEGRATION.md section 1 and runs it
UNCHANGED against the GPU path.  Run here:  python tests/golden/make_worker_test_fixture.py"""
import base64
import hashlib
import os
import zlib

SRC = "/root/reference/worker_test.py"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "worker_test.fixture")

if __name__ == "__main__":
    raw = open(SRC, "rb").read()
    body = base64.b85encode(zlib.compress(raw, 9)).decode()
    with open(DST, "w") as f:
        f.write("# vainglorygame/analyzer worker_test.py (unmodified), zlib + base85; made by make_worker_test_fixture.py\n")
        f.write("# sha256 %s bytes %d\n" % (hashlib.sha256(raw).hexdigest(), len(raw)))
        for i in range(0, len(body), 100):
            f.write(body[i:i + 100] + "\n")
    print("wrote", DST, len(raw), "->", os.path.getsize(DST))
