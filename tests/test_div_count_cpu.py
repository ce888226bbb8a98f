"""The GPU's AGG block mean divides by a compile-time count with q0 = a*rn(1/N), r = fma(-N, q0, a), q = fma(r, rn(1/N), q0)
instead of the general IEEE division (csrc/srnn_device.cuh: div_count).  The arithmetic is restated in C
(tests/div_count_check.c) and compared with `/` bit for bit: N = 1..64, random doubles over the whole exponent range,
sums of float32 values as the kernel forms them, infinities included."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def test_constant_count_division_is_correctly_rounded(tmp_path):
    exe = str(tmp_path / "div_count_check")
    subprocess.check_call(["/usr/bin/gcc", "-O2", "-march=x86-64-v3", "-ffp-contract=off", "-o", exe,
                           os.path.join(HERE, "div_count_check.c"), "-lm"])
    out = subprocess.run([exe, "200000"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:]
    last = out.stdout.strip().splitlines()[-1].split()
    assert last[0] == "checked" and int(last[1]) > 10_000_000 and int(last[3]) == 0, out.stdout[-500:]
