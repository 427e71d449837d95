"""The multi-GPU entry points (la_comm_create, la_*_sharded) from a plain C++ program: tests/comm_two_proc.cpp, one process
per GPU, exchange callback = files in a scratch directory.  No torch, no NCCL, no MPI on that path -- what a Rust or C++ user
of liblinalg_b200.so has.  Compiles everywhere; runs where at least 2 GPUs are visible (gpurun --gpus 2)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build(tmp_path):
    exe = str(tmp_path / "comm_two_proc")
    libdir = os.path.join(ROOT, "linear_algebra_b200")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", os.path.join(ROOT, "tests", "comm_two_proc.cpp"), "-o", exe,
                    "-L" + libdir, "-llinalg_b200", "-Wl,-rpath," + libdir, "-lpthread"], check=True)
    return exe


def test_comm_program_compiles(tmp_path):
    build(tmp_path)


@pytest.mark.gpu
@pytest.mark.parametrize("peer_memory", [True, False])
def test_sharded_entry_points_from_cpp_processes(tmp_path, peer_memory):
    """peer_memory False: LA_COMM_NO_PEER_MEMORY makes la_comm_create behave as on a box without peer access, so the fallbacks
    through the exchange callback (and the refusal of the sharded transpose) are exercised too."""
    import ctypes
    from linear_algebra_b200 import _lib
    n = ctypes.c_int(0)
    _lib.call("la_device_count", ctypes.byref(n))
    if n.value < 2:
        pytest.skip("needs >= 2 GPUs (run under gpurun --gpus 2)")
    world = 2 if n.value < 4 else 4
    exe = build(tmp_path)
    scratch = tmp_path / "xchg"; scratch.mkdir()
    env = dict(os.environ)
    env.pop("LA_COMM_NO_PEER_MEMORY", None)
    if not peer_memory:
        env["LA_COMM_NO_PEER_MEMORY"] = "1"
    procs = [subprocess.Popen([exe, str(r), str(world), str(scratch)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env) for r in range(world)]
    outs = [p.communicate(timeout=300) for p in procs]
    for r, (p, (out, err)) in enumerate(zip(procs, outs)):
        assert p.returncode == 0, "rank %d: %s %s" % (r, out[-2000:], err[-2000:])
        assert "rank %d/%d ok" % (r, world) in out
