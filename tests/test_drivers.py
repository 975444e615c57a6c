"""The C++ drop-in drivers (drivers/{single,weak,strong,weak4d}.cpp): they build from the host headers and
the C ABI alone, refuse to run without a GPU, and on a GPU validate themselves (-v) and print the
reference's result lines (weak/main.cu:304-318, strong/main.cu:550-566)."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "drivers", "_build")


@pytest.fixture(scope="module")
def drivers(built):
    have = all(os.path.exists(os.path.join(BIN, n)) for n in ("single", "weak", "strong", "weak4d"))
    if not have or os.path.exists("/usr/bin/g++"):
        subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "drivers")])
    return BIN


def _run(path, *args, timeout=600):
    p = subprocess.run([path] + list(args), capture_output=True, text=True, timeout=timeout)
    return p.returncode, p.stdout + p.stderr


def test_drivers_build_without_cuda_headers(drivers):
    for n in ("single", "weak", "strong", "weak4d"):
        src = open(os.path.join(ROOT, "drivers", n + ".cpp")).read()
        assert "cuda_runtime" not in src and "<mpi.h>" not in src
        assert os.access(os.path.join(drivers, n), os.X_OK)
    rc, out = _run(os.path.join(drivers, "weak"), "-h")
    assert rc == 0 and "-s: comma separated Int[3], per-process domain size" in out


def test_drivers_refuse_to_run_without_a_gpu(drivers, built):
    import ctypes
    n = ctypes.c_int()
    if built.lib().bricks_device_count(ctypes.byref(n)) == 0 and n.value > 0:
        pytest.skip("a GPU is present")
    for n, args in (("single", ["-d", "32"]), ("weak", ["-s", "32", "-I", "1"]), ("strong", ["-d", "64", "-s", "32", "-I", "1"]),
                    ("weak4d", ["-s", "16", "-I", "1"])):
        rc, out = _run(os.path.join(drivers, n), *args)
        assert rc != 0 and "CUDA device" in out, (n, rc, out)


def _perf(out):
    m = re.search(r"^perf ([0-9.eE+-]+) GStencil/s", out, re.M)
    assert m, out
    return float(m.group(1))


@pytest.mark.gpu
def test_single_driver_validates(drivers):
    rc, out = _run(os.path.join(drivers, "single"), "-d", "64", "-I", "3")
    assert rc == 0, out
    assert "d3pt7" in out and "cond" in out and "d3pt27" in out and "Arr: " in out and "Bri: " in out and "mismatch" not in out
    for w in (7, 13, 25, 125):
        assert "mpi%dpt" % w in out


@pytest.mark.gpu
@pytest.mark.parametrize("which", (7, 13, 25, 125))
def test_weak_driver_validates(drivers, which):
    rc, out = _run(os.path.join(drivers, "weak"), "-s", "64,48,32", "-I", "3", "-S", str(which), "-v")
    assert rc == 0, out
    for key in ("Bri: ", "calc ", "move ", "call ", "wait ", "  | MPI size (MB): ", "  | MPI speed (GB/s): "):
        assert key in out, out
    assert _perf(out) > 0 and "mismatch" not in out
    # -v compares with the ARRAY twin on the host (weak/main.cu:143-214, 321-323), not with another kernel
    m = re.search(r"result match: max relative error vs the host array twin ([0-9.eE+-]+)", out)
    assert m and float(m.group(1)) <= 1e-12, out


@pytest.mark.gpu
@pytest.mark.parametrize("flag", ("-m", "-p", "-n"))
def test_weak_driver_validates_with_the_other_exchanges(drivers, flag):
    """-m: MultiStageExchangeView, -p: one packed message per peer, -n: the serial schedule; the host
    array twin exchanges periodically on its own, so every form must agree with it"""
    rc, out = _run(os.path.join(drivers, "weak"), "-s", "64,48,32", "-I", "3", "-S", "13", "-v", flag)
    assert rc == 0, out
    assert _perf(out) > 0 and "mismatch" not in out and "result match" in out


@pytest.mark.gpu
@pytest.mark.parametrize("args", (["-s", "32"], ["-s", "16,24,16,20", "-a"], ["-s", "20,20,28,16", "-t"]))
def test_weak4d_driver_validates(drivers, args):
    """the 4-D program (weak/main-4d.cpp): exchange + 8 shrinking sweeps against the host array twin; -a: the
    page-aligned numbering, -t: the twin kernel"""
    rc, out = _run(os.path.join(drivers, "weak4d"), *args, "-I", "3", "-v")
    assert rc == 0, out
    for key in ("d4pt9 on B200", "Bri: ", "calc ", "call ", "wait ", "  | MPI size (MB): ", "Total of 312 parts"):
        assert key in out, out
    assert _perf(out) > 0 and "mismatch" not in out
    m = re.search(r"result match: max relative error vs the host array twin ([0-9.eE+-]+)", out)
    assert m and float(m.group(1)) <= 1e-12, out


@pytest.mark.gpu
@pytest.mark.parametrize("which", (7, 125))
def test_strong_driver_validates(drivers, which):
    rc, out = _run(os.path.join(drivers, "strong"), "-d", "128", "-s", "32", "-I", "3", "-S", str(which), "-v")
    assert rc == 0, out
    for key in ("Bri: ", "calc : ", "pack : ", "call : ", "wait : ", "move : ", "part "):
        assert key in out, out
    assert _perf(out) > 0 and "Validated" in out and "mismatch" not in out
    m = re.search(r"Validated, max relative error vs the global periodic array ([0-9.eE+-]+)", out)
    assert m and float(m.group(1)) <= 1e-12, out


@pytest.mark.gpu
@pytest.mark.parametrize("ngpus", (2, 8))
def test_drivers_multi_gpu_validate(drivers, built, ngpus):
    import ctypes
    n = ctypes.c_int()
    built.check(built.lib().bricks_device_count(ctypes.byref(n)))
    if n.value < ngpus:
        pytest.skip("needs %d GPUs" % ngpus)
    rc, out = _run(os.path.join(drivers, "weak"), "-G", str(ngpus), "-s", "64", "-I", "3", "-S", "125", "-v")
    assert rc == 0 and "mismatch" not in out, out
    rc, out = _run(os.path.join(drivers, "weak"), "-G", str(ngpus), "-s", "64", "-I", "3", "-S", "7", "-v", "-m")
    assert rc == 0 and "mismatch" not in out, out
    rc, out = _run(os.path.join(drivers, "weak"), "-G", str(ngpus), "-s", "64", "-I", "3", "-S", "13", "-v", "-p")
    assert rc == 0 and "mismatch" not in out, out
    rc, out = _run(os.path.join(drivers, "weak"), "-G", str(ngpus), "-s", "64", "-I", "3", "-S", "25", "-v", "-E")
    assert rc == 0 and "mismatch" not in out, out
    rc, out = _run(os.path.join(drivers, "weak"), "-G", str(ngpus), "-s", "64", "-I", "3", "-S", "125", "-v", "-N")
    assert rc == 0 and "mismatch" not in out and "peer-mapped" not in out, out
    rc, out = _run(os.path.join(drivers, "weak"), "-G", str(ngpus), "-s", "64", "-I", "3", "-S", "7", "-v", "-e")
    assert rc == 0 and "mismatch" not in out and "peer-mapped" in out, out
    rc, out = _run(os.path.join(drivers, "strong"), "-G", str(ngpus), "-d", "128", "-s", "32", "-I", "3", "-v")
    assert rc == 0 and "Validated" in out, out
