"""The C++ host mirror (brpc-github-actions_b200/host) over libb2c.so: compile it with g++ and,
on a GPU box, run the reference's own unit tests re-pointed at it (tests/cpp/test_host_api.cpp)."""
import os
import shutil
import subprocess

import pytest

import brpc_b200
from oracle import oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "brpc-github-actions_b200")
BIN = os.path.join(ROOT, "tests", "cpp", "test_host_api.bin")


def _build():
    O.port()  # makes sure oracle/libb2oracle.so exists
    brpc_b200.load()
    cxx = shutil.which("g++")
    if cxx is None:
        if os.path.exists(BIN):
            return BIN
        pytest.skip("no g++ and no prebuilt test binary")
    cmd = [cxx, "-std=c++17", "-O1", "-Wall", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(PKG, "host"),
           "-I", os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests", "cpp", "test_host_api.cpp"),
           os.path.join(PKG, "host", "b200_codec.cpp"),
           os.path.join(PKG, "host", "b200_framing.cpp"), os.path.join(PKG, "host", "b200_block_pool.cpp"), "-L", PKG, "-lb2c", "-L", os.path.join(ROOT, "oracle"),
           "-lb2oracle", f"-Wl,-rpath,{PKG}", f"-Wl,-rpath,{os.path.join(ROOT, 'oracle')}", "-lpthread", "-o", BIN]
    subprocess.run(cmd, check=True)
    return BIN


SHIM_BIN = os.path.join(ROOT, "tests", "cpp", "test_brpc_shim.bin")


def _build_shim():
    """host/brpc_gpu_snappy.cpp (the COMPRESS_TYPE_SNAPPY handler bodies) against the stand-in bRPC headers of
    tests/cpp/mock_brpc/ (butil::IOBuf, protobuf Message: absent from this image)."""
    O.port()
    brpc_b200.load()
    cxx = shutil.which("g++")
    if cxx is None:
        if os.path.exists(SHIM_BIN):
            return SHIM_BIN
        pytest.skip("no g++ and no prebuilt test binary")
    cmd = [cxx, "-std=c++17", "-O1", "-Wall", "-DB200_WITH_BRPC", "-I", os.path.join(ROOT, "include"),
           "-I", os.path.join(PKG, "host"), "-I", os.path.join(ROOT, "oracle"), "-I", os.path.join(ROOT, "tests", "cpp", "mock_brpc"),
           os.path.join(ROOT, "tests", "cpp", "test_brpc_shim.cpp"), os.path.join(PKG, "host", "brpc_gpu_snappy.cpp"),
           os.path.join(PKG, "host", "b200_codec.cpp"), os.path.join(PKG, "host", "b200_block_pool.cpp"),
           "-L", PKG, "-lb2c", "-L", os.path.join(ROOT, "oracle"), "-lb2oracle", f"-Wl,-rpath,{PKG}",
           f"-Wl,-rpath,{os.path.join(ROOT, 'oracle')}", "-lpthread", "-o", SHIM_BIN]
    subprocess.run(cmd, check=True)
    return SHIM_BIN


def test_host_mirror_compiles_and_links():
    assert os.path.exists(_build())


def test_brpc_shim_compiles_against_stand_in_headers():
    assert os.path.exists(_build_shim())


@pytest.mark.gpu
def test_brpc_shim_handlers_from_32_threads():
    r = subprocess.run([_build_shim()], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all checks passed" in r.stdout


@pytest.mark.gpu
def test_reference_unit_tests_on_the_gpu_codec():
    r = subprocess.run([_build()], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all checks passed" in r.stdout
