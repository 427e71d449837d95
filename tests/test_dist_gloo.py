"""World-size-2 gloo tests (CPU) of the multi-GPU bookkeeping in linear_algebra_b200/dist.py: row-block partitioning
(la_partition), which shard goes where, the byte all-gather that la_comm_create takes as its exchange callback.  The per-rank
arithmetic and the rank-ordered combinations are a numpy stand-in (tests/fake_local.py: Vector / Matrix / ShardedEngine); on
GPUs the same bookkeeping drives liblinalg_b200's la_*_sharded entry points (tests/test_gpu_dist.py, tests/test_gpu_comm_cpp.py)."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def test_partition_covers_everything_once():
    from linear_algebra_b200.dist import partition
    for n in (0, 1, 7, 8, 9, 1000, 8192):
        for world in (1, 2, 3, 8):
            spans = [partition(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def test_the_host_model_combines_in_rank_order_with_the_element_types_arithmetic():
    sys.path.insert(0, HERE)
    from fake_local import combine_scalars
    parts = [np.uint8(200).tobytes(), np.uint8(100).tobytes()]
    assert combine_scalars(parts, np.uint8) == np.uint8(44)
    f = [np.float32(1e8).tobytes(), np.float32(1.0).tobytes(), np.float32(-1e8).tobytes()]
    assert combine_scalars(f, np.float32) == np.float32(np.float32(np.float32(1e8) + np.float32(1.0)) + np.float32(-1e8))


def _worker(rank, world, port, q):
    try:
        sys.path.insert(0, HERE); sys.path.insert(0, os.path.dirname(HERE))
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
        import torch.distributed as td
        td.init_process_group("gloo", rank=rank, world_size=world)
        import fake_local
        from linear_algebra_b200 import dist
        comm = dist.Comm()
        rng = np.random.default_rng(7)                # same data on every rank
        for np_dt in (np.int32, np.uint8, np.float32):
            n = 1001
            a = rng.integers(0, 200, n).astype(np_dt); b = rng.integers(1, 200, n).astype(np_dt)
            A = dist.ShardedVector.from_global(comm, a, local_mod=fake_local); B = dist.ShardedVector.from_global(comm, b, local_mod=fake_local)
            assert A.local.len() == dist.partition(n, world, rank)[1] - dist.partition(n, world, rank)[0]
            with np.errstate(over="ignore"):
                assert np.array_equal((A + B).gather(), (a + b).astype(np_dt))
                assert np.array_equal((A - B).gather(), (a - b).astype(np_dt))
                assert np.array_equal((A * B).gather(), (a * b).astype(np_dt))
                assert np.array_equal((A * np_dt(3)).gather(), (a * np_dt(3)).astype(np_dt))
                if np.dtype(np_dt).kind != "f":
                    assert A.sum() == a.sum(dtype=np_dt)
                    assert dist.dot(A, B) == (a * b).sum(dtype=np_dt)
                else:
                    assert abs(float(A.sum()) - float(a.sum(dtype=np.float64))) < 1e-2
            assert A == A and not (A == B)
            C = dist.ShardedVector.from_global(comm, a, local_mod=fake_local)
            C += B; C -= B
            assert C == A
            # matrices: 37 x 12, rows sharded
            rows, cols = 37, 12
            m = rng.integers(0, 20, rows * cols).astype(np_dt); v_r = rng.integers(0, 20, rows).astype(np_dt); v_c = rng.integers(0, 20, cols).astype(np_dt)
            M = dist.ShardedMatrix.from_global(comm, m, rows, cols, local_mod=fake_local)
            with np.errstate(over="ignore"):
                expect_mv = (m.reshape(rows, cols).astype(np.int64) @ v_c.astype(np.int64)).astype(np_dt)
                expect_vm = (v_r.astype(np.int64) @ m.reshape(rows, cols).astype(np.int64)).astype(np_dt)
            got_mv = dist.mul_transpose_mat(fake_local.Vector.from_vec(v_c), M)
            assert np.array_equal(got_mv.gather(), expect_mv)
            got_vm = dist.ShardedVector.from_global(comm, v_r, local_mod=fake_local) * M          # replicated result
            assert np.array_equal(got_vm.to_vec(), expect_vm)
            Bm = fake_local.Matrix.from_vec(rng.integers(0, 20, cols * 5).astype(np_dt), cols, 5)
            Cm = M * Bm
            with np.errstate(over="ignore"):
                expect_mm = (m.reshape(rows, cols).astype(np.int64) @ Bm._m().astype(np.int64)).astype(np_dt)
            assert np.array_equal(Cm.gather().reshape(rows, 5), expect_mm)
            assert np.array_equal((M + M).gather(), (m + m).astype(np_dt))
        comm.barrier()
        td.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "FAIL: " + traceback.format_exc()))


@pytest.mark.timeout(300)
def test_sharded_ops_world_size_2_gloo():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(msg == "ok" for _, msg in results), results
