"""CPU-side checks of the product library: it builds, loads, and exports every symbol that
include/orbx.h declares; without a GPU every compute entry fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def so(pkg):
    return pkg.orbx.build()


def test_exports_every_declared_symbol(so):
    hdr = open(os.path.join(ROOT, "include", "orbx.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(orbx_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 20
    lib = C.CDLL(so)
    missing = [n for n in sorted(names) if not hasattr(lib, n)]
    assert not missing, missing


def test_no_oracle_in_product():
    """The product path must not link or import the oracle."""
    out = subprocess.run(["ldd", os.path.join(ROOT, "orb_slam3_v1.0_b200", "lib", "liborbx.so")],
                         capture_output=True, text=True).stdout
    assert "oracle" not in out
    for dp, _, fs in os.walk(os.path.join(ROOT, "orb_slam3_v1.0_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".h", ".cuh")):
                txt = open(os.path.join(dp, f)).read()
                assert "orb_oracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f


def test_fails_loudly_without_gpu(pkg, so):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    ex = pkg.orbx.ORBextractor()
    with pytest.raises(pkg.orbx.OrbxError):
        ex.prepare(752, 480, 1)


def test_stdsort_emulation_matches_libstdcxx(tmp_path):
    exe = str(tmp_path / "t")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe, os.path.join(ROOT, "tests", "cpp", "test_stdsort.cpp")])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.startswith("OK"), out.stdout


def build_shim_test(so, out_dir):
    """Compiles tests/cpp/test_shim.cpp (C++ drop-in headers include/orbx/*.h) against liborbx.so."""
    exe = os.path.join(str(out_dir), "test_shim")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-o", exe,
                           os.path.join(ROOT, "tests", "cpp", "test_shim.cpp"), "-L" + os.path.dirname(so), "-lorbx",
                           "-Wl,-rpath," + os.path.dirname(so)])
    return exe


def test_cpp_shim_compiles_and_fails_loudly_without_gpu(so, tmp_path):
    exe = build_shim_test(so, tmp_path)
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    img = tmp_path / "img.raw"
    img.write_bytes(bytes(128 * 96))
    r = subprocess.run([exe, str(img), "128", "96", "100", str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 1 and "no CUDA device" in r.stderr      # the shim throws; no CPU path


def test_octree_sort_closed_form_matches_std_sort(tmp_path):
    """extractor.cu evaluates std::sort's Hoare partition in closed form per warp and replaces the
    final insertion sort by a stable rank sort; the same formulation must equal libstdc++'s
    std::sort on tie-heavy (size, UL.x) keys (reference src/ORBextractor.cc:209-224, :373)."""
    exe = str(tmp_path / "tpp")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-w", "-o", exe,
                           os.path.join(ROOT, "tests", "cpp", "test_parallel_partition.cpp")])
    r = subprocess.run([exe, "20000"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and r.stdout.startswith("OK"), r.stdout + r.stderr


def test_reference_signature_layer_and_opencv_overload_compile(so, tmp_path):
    """include/orbx/ORBmatcher_ref.h (the reference's static ORBmatcher signatures as templates, instantiated with the
    mock SLAM types) and the ORBX_WITH_OPENCV overload of ORBextractor (against tests/cpp/opencv_stub) compile
    warning-free; they are RUN by tests/test_gpu_cpp_api.py on the GPU."""
    cpp = os.path.join(ROOT, "tests", "cpp")
    link = ["-L" + os.path.dirname(so), "-lorbx", "-Wl,-rpath," + os.path.dirname(so)]
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-o", str(tmp_path / "a"), os.path.join(cpp, "test_ref_api.cpp")] + link)
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-DORBX_WITH_OPENCV", "-I" + os.path.join(cpp, "opencv_stub"), "-pthread",
                           "-o", str(tmp_path / "b"), os.path.join(cpp, "test_shim_opencv.cpp")] + link)
