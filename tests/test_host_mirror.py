"""C++ host mirror of the RLGC plugin surface (reinforcement-learning_b200/host/RLGymCPP/EnvSet.h) over the C ABI.

CPU: the mirror compiles against include/gigaenv.h, links libgigaenv_cuda.so, translates the stock plugin objects into the
device descriptor and refuses a user subclass loudly (no device call is made). GPU: the ExampleMain-style loop
(EnvCreateFn -> EnvSet -> Reset / StepFirstHalf(async) / Sync / StepSecondHalf) runs on cuda:0 through the mirror."""
import importlib
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "reinforcement-learning_b200")
SRC = os.path.join(ROOT, "tests", "host", "test_envset_host.cpp")
OUT = os.path.join(ROOT, "tests", "host", "_build", "test_envset_host")


def build_host_test():
    """g++ only (no nvcc, no CUDA headers): the mirror sees the device through the C ABI alone."""
    hdr = os.path.join(PKG, "host", "RLGymCPP", "EnvSet.h")
    lib = os.path.join(PKG, "libgigaenv_cuda.so")
    deps = [SRC, hdr, os.path.join(ROOT, "include", "gigaenv.h"), lib]
    if os.path.exists(OUT) and all(os.path.getmtime(OUT) >= os.path.getmtime(d) for d in deps):
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(PKG, "host"), SRC, "-o", OUT,
                           "-L", PKG, "-lgigaenv_cuda", "-Wl,-rpath," + PKG, "-Wl,-rpath,/usr/local/cuda/lib64"])
    return OUT


def build_experience_test():
    src = os.path.join(ROOT, "tests", "host", "test_experience_host.cpp")
    out = os.path.join(ROOT, "tests", "host", "_build", "test_experience_host")
    deps = [src, os.path.join(PKG, "host", "GigaLearnCPP", "Experience.h"), os.path.join(ROOT, "include", "gigaenv.h"), os.path.join(PKG, "libgigaenv_cuda.so")]
    if os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in deps):
        return out
    os.makedirs(os.path.dirname(out), exist_ok=True)
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-I", os.path.join(PKG, "host"), src, "-o", out,
                           "-L", PKG, "-lgigaenv_cuda", "-Wl,-rpath," + PKG, "-Wl,-rpath,/usr/local/cuda/lib64"])
    return out


def test_learner_side_mirror_compiles_without_cuda_headers(ggl):
    ggl.lib()
    assert os.path.exists(build_experience_test())


@pytest.mark.gpu
def test_learner_side_mirror_on_gpu(ggl):
    """GGL::ExperienceBuffer (two epochs against std::shuffle of the same engine and seed) and GGL::TrajectoryStore (the
    hand-checked order of tests/test_traj_oracle_cpu.py) through the C++ mirror."""
    exe = build_experience_test()
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "experience ok=1" in r.stdout and "trajectory ok=1 bad_desc_throws=1" in r.stdout


def test_mirror_translates_plugins_and_rejects_user_subclasses(ggl):
    ggl.lib()
    exe = build_host_test()
    r = subprocess.run([exe, "describe"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "ok=1 user_subclass_throws=1 roster_check=1" in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("ppt", [1, 2])
def test_mirror_runs_example_loop_on_gpu(ggl, tmp_path, ppt):
    mesh = importlib.import_module("reinforcement-learning_b200.arena_mesh")
    d = tmp_path / "collision_meshes" / "soccar"
    d.mkdir(parents=True)
    for i, b in enumerate(mesh.soccar_meshes()):
        (d / ("mesh_%02d.cmf" % i)).write_bytes(b)
    exe = build_host_test()
    r = subprocess.run([exe, "run", str(tmp_path / "collision_meshes"), str(ppt), "64", "12"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "run ok players=%d" % (64 * 2 * ppt) in r.stdout and "finite=1" in r.stdout and "size_check=1" in r.stdout
