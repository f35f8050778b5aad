"""Tests of the C++ host side (stencil_b200/host): the reference's CLI surface
(src/program_options.cpp:13-37), its two report lines (src/main.cpp:33-43) that
run_expr.py scrapes (run_expr.py:9), exit codes (main.cpp:59-61) and the -c self-check."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "stencil_b200", "host", "build", "bin", "stencil_main")
AVG = r'The average time taken by (.*) method is (.*)ms for \d+ iterations\.'   # run_expr.py:9


def run(*args):
    return subprocess.run([BIN, *map(str, args)], capture_output=True, text=True, timeout=600)


def test_cpu_methods_report_lines_and_check(built):
    r = run("-s", 64, "-i", 10, "-b", 8, "-R", 2, "-m", "CPU", "CPUOMP", "-c")
    assert r.returncode == 0, r.stderr
    out = r.stdout
    for m in ("CPU", "CPUOMP"):
        assert f"Start to check the correctness of method {m}." in out
        assert f"The results of method {m} is correct." in out
        assert len(re.findall(rf"^{m} Method spent .*ms for 10 iterations\.$", out, re.M)) == 2
    assert [m for m, _ in re.findall(AVG, out)] == ["CPU", "CPUOMP"]


def test_required_options_and_exit_code(built):
    assert run("-s", 64).returncode == 1                       # -i, -b, -m missing
    assert run("-s", 64, "-i", 1, "-b", 8).returncode == 1     # -m missing
    assert run("-s", 64, "-i", 1, "-b", 8, "-m", "CPU", "--bogus").returncode == 1
    assert run("-s", "abc", "-i", 1, "-b", 8, "-m", "CPU").returncode == 1
    assert run("--help").returncode == 0


def test_long_names_defaults_and_extensions(built):
    r = run("--matrix-size", 32, "--iteration", 3, "--block-size", 4, "--radius", 2, "--repeat", 1,
            "--methods", "CPU", "--check-result", "--dtype", "f64", "--stencil", "box", "--init", "random")
    assert r.returncode == 0 and "is correct" in r.stdout
    r = run("-s", 1, "-i", 2, "-b", 1, "-m", "CPU", "--shape", "9,8,7", "--dtype=f64", "-c")
    assert r.returncode == 0 and "is correct" in r.stdout
    assert run("-s", 8, "-i", 1, "-b", 1, "-m", "CPU", "--dtype", "f16").returncode == 1


def test_legacy_and_unknown_methods_fail_loudly(built):
    r = run("-s", 16, "-i", 1, "-b", 2, "-m", "RMA")
    assert r.returncode == 3 and "not available on the B200 backend" in r.stderr
    r = run("-s", 16, "-i", 1, "-b", 2, "-m", "Nope")
    assert r.returncode == 3 and "unknown method name" in r.stderr


def test_cuda_method_without_gpu_is_an_error_not_a_fallback(built):
    from stencil_b200 import capi
    if capi.lib().sb200_device_count() > 0:
        pytest.skip("a GPU is visible")
    r = run("-s", 16, "-i", 1, "-b", 2, "-m", "CUDA")
    assert r.returncode == 3 and "no CPU fallback" in r.stderr
    assert "average time" not in r.stdout


def test_bmp_dump(built, tmp_path):
    p = tmp_path / "f.bmp"
    r = run("-s", 20, "-i", 30, "-b", 4, "-m", "CPU", "--bmp", p)
    assert r.returncode == 0
    data = p.read_bytes()
    assert data[:2] == b"BM" and len(data) == 54 + 20 * 60
    # byte-identical to the file the reference's own Stencil::to_bmp + BMPImage::save write for
    # the same field (tests/golden/make_golden.py drives them through oracle/_ref/libref.so)
    gold = open(os.path.join(ROOT, "tests", "golden", "ref_bmp_N20_T30.bmp"), "rb").read()
    assert data == gold


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [
    ["-s", 4096, "-i", 100, "-b", 64],                                         # config 1 as the reference CLI runs it (f32)
    ["-s", 400, "-i", 1000, "-b", 50],                                         # reference run.sh:3
    ["-s", 512, "-i", 20, "-b", 64, "--dtype", "f64", "--init", "random"],
    ["-s", 1, "-i", 6, "-b", 1, "--shape", "160,120,90", "--dtype", "f64", "--init", "random"],
    ["-s", 300, "-i", 12, "-b", 1, "--stencil", "box", "--init", "random"],
    ["-s", 200, "-i", 5, "-b", 1, "-r", 3],
    ["-s", 1, "-i", 3, "-b", 1, "--shape", "130,70,40", "-r", 2, "--stencil", "box", "--dtype", "f64", "--init", "random"],  # 125-point
    ["-s", 1, "-i", 37, "-b", 1, "--shape", "100003", "-r", 2, "--dtype", "f64", "--init", "random"],  # 1-D
    ["-s", 1, "-i", 10, "-b", 1, "--shape", "4096", "--stencil", "box", "--kernel", "line1d"],
])
def test_cuda_method_self_check(gpu, extra):
    r = run(*extra, "-m", "CUDA", "-c")
    assert r.returncode == 0, r.stdout + r.stderr
    assert "The results of method CUDA is correct." in r.stdout
    assert "bit-exact vs naive loop: yes" in r.stdout
    assert re.search(AVG, r.stdout)


@pytest.mark.gpu
@pytest.mark.parametrize("extra", [
    ["-s", 700, "-i", 24, "-b", 1, "--stencil", "box", "--init", "random"],                                   # 9-point, fp32
    ["-s", 1, "-i", 8, "-b", 1, "--shape", "130,120,70", "--stencil", "box", "--dtype", "f64", "--init", "random"],  # 27-point
])
def test_cuda_reassociate_passes_the_reference_check_by_tolerance(gpu, extra):
    """--reassociate (SB200_FLAG_REASSOCIATE): the reference's own -c check is a tolerance check
    (stencil.cpp:75-151, 1e-4 absolute), which the separable sums pass without being bit-exact."""
    r = run(*extra, "-m", "CUDA", "-c", "--reassociate")
    assert r.returncode == 0, r.stdout + r.stderr
    assert "The results of method CUDA is correct." in r.stdout
    assert "bit-exact vs naive loop: no" in r.stdout


def test_reassociate_with_resident_check_is_refused(built):
    r = run("-s", 64, "-i", 2, "-b", 1, "-m", "CUDA", "-c", "--resident", "--reassociate")
    assert r.returncode != 0 and "naive order" in (r.stdout + r.stderr)
