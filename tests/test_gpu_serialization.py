"""Section 8 row f3: the reference's file harness (src/sort.h:15-103 + src/main.cpp:9-44, compiled unmodified into
oracle/_ref/ref_main) over the shim's Serial:: layer.  tests/native/fherma_roundtrip.cpp plays the key owner: it serialises
context, public key, relinearisation key, rotation keys and an encrypted array, runs the reference's main as a separate
process without the secret key, deserialises the output ciphertext and decrypts it (max error < 0.01, tests/DirectSortTest.cpp:169).
The wire format is this engine's own raw-limb container, not OpenFHE's cereal archives (unverifiable without OpenFHE)."""
import os
import subprocess
import tempfile

import pytest

pytestmark = pytest.mark.gpu

REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref")


@pytest.mark.parametrize("logn", [13])
def test_reference_main_round_trip(logn):
    drv, main = os.path.join(REF, "fherma_roundtrip"), os.path.join(REF, "ref_main")
    if not (os.path.exists(drv) and os.path.exists(main)):
        pytest.skip("oracle/_ref/ref_main not built (no reference checkout on the build machine)")
    with tempfile.TemporaryDirectory() as d:
        r = subprocess.run([drv, main, d, str(logn)], capture_output=True, text=True, timeout=900)
        print(r.stdout[-500:], r.stderr[-2000:])
        assert r.returncode == 0, r.stderr[-2000:]
        assert "max_abs_err" in r.stdout
        sizes = {f: os.path.getsize(os.path.join(d, f)) for f in os.listdir(d)}
        assert set(sizes) == {"cc.bin", "pk.bin", "mult.bin", "rot.bin", "in.bin", "out.bin"}
        assert sizes["cc.bin"] == 64 and sizes["rot.bin"] > 10 * sizes["mult.bin"]
