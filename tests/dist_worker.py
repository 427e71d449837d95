"""Worker for tests/test_gpu_dist.py: run under torchrun, one rank per GPU, NCCL."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, HERE)


def main():
    import torch
    import torch.distributed as td
    import linear_algebra_b200 as la
    from linear_algebra_b200 import dist
    import oracle
    rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    td.init_process_group("nccl", device_id=torch.device("cuda", local))
    la.init(local)
    comm = dist.Comm()
    P = oracle.port()
    rng = np.random.default_rng(11)                   # same data on every rank
    for np_dt in (np.int32, np.uint16, np.float32):
        n = (1 << 20) + 77
        if np.dtype(np_dt).kind == "f":
            a = rng.uniform(-1, 1, n).astype(np_dt); b = rng.uniform(-1, 1, n).astype(np_dt)
        else:
            info = np.iinfo(np_dt)
            a = rng.integers(info.min, info.max, n, dtype=np_dt, endpoint=True); b = rng.integers(info.min, info.max, n, dtype=np_dt, endpoint=True)
        A = dist.ShardedVector.from_global(comm, a); B = dist.ShardedVector.from_global(comm, b)
        assert (A + B).gather().tobytes() == P.ew_vec_vec(oracle.ADD, a, b).tobytes()
        assert (A * B).gather().tobytes() == P.ew_vec_vec(oracle.MUL, a, b).tobytes()
        assert A == A and not (A == B)
        if np.dtype(np_dt).kind != "f":
            assert A.sum() == P.sum_vec(a) and dist.dot(A, B) == P.dot_vec_vec(a, b)
        else:
            d64, ad = P.f32_dot_truth(a, b)
            assert abs(float(dist.dot(A, B)) - d64) <= 2.0 ** -22 * ad
        rows, cols = 1037, 2048
        if np.dtype(np_dt).kind == "f":
            m = rng.uniform(-1, 1, rows * cols).astype(np_dt); vr = rng.uniform(-1, 1, rows).astype(np_dt); vc = rng.uniform(-1, 1, cols).astype(np_dt)
            bm = rng.uniform(-1, 1, cols * 256).astype(np_dt)
        else:
            m = rng.integers(info.min, info.max, rows * cols, dtype=np_dt, endpoint=True); vr = rng.integers(info.min, info.max, rows, dtype=np_dt, endpoint=True)
            vc = rng.integers(info.min, info.max, cols, dtype=np_dt, endpoint=True); bm = rng.integers(info.min, info.max, cols * 256, dtype=np_dt, endpoint=True)
        M = dist.ShardedMatrix.from_global(comm, m, rows, cols)
        mv = dist.mul_transpose_mat(la.Vector.from_vec(vc), M).gather()
        vm = (dist.ShardedVector.from_global(comm, vr) * M).to_vec()
        mm = (M * la.Matrix.from_vec(bm, cols, 256)).gather()
        if np.dtype(np_dt).kind != "f":
            assert mv.tobytes() == P.mul_vec_transpose_mat(vc, m, rows, cols).tobytes()
            assert vm.tobytes() == P.mul_vec_mat(vr, m, rows, cols).tobytes()
            assert mm.tobytes() == P.mul_mat_mat(m, bm, rows, cols, 256).tobytes()
        else:
            for got, ref in ((mv, P.mul_vec_transpose_mat(vc, m, rows, cols)), (vm, P.mul_vec_mat(vr, m, rows, cols)), (mm, P.mul_mat_mat(m, bm, rows, cols, 256))):
                assert np.linalg.norm(got - ref) / np.linalg.norm(ref) <= 1e-5
        # sharded transpose (all_to_all over NVLink) and the NVLink-gathered host GEMM
        T = dist.transpose(M)
        assert T.len() == (cols, rows)
        assert T.gather().tobytes() == P.transpose(m, rows, cols).tobytes()
        Mh, Kh, Nh = 1024, 512 * world, 384
        if np.dtype(np_dt).kind == "f":
            ah = rng.uniform(-1, 1, Mh * Kh).astype(np_dt); bh = rng.uniform(-1, 1, Kh * Nh).astype(np_dt)
        else:
            ah = rng.integers(info.min, info.max, Mh * Kh, dtype=np_dt, endpoint=True); bh = rng.integers(info.min, info.max, Kh * Nh, dtype=np_dt, endpoint=True)
        lo, hi = dist.partition(Mh, world, rank)
        ch = dist.mul_host(comm, ah[lo * Kh:hi * Kh], bh, Mh, Kh, Nh)
        ref = P.mul_mat_mat(ah, bh, Mh, Kh, Nh)[lo * Nh:hi * Nh]
        if np.dtype(np_dt).kind != "f":
            assert ch.tobytes() == ref.tobytes()
        else:
            assert np.linalg.norm(ch - ref) / np.linalg.norm(ref) <= 1e-5
    # the fused reduce + NVLink exchange must be the path in use, and stay consistent over many rounds (parity reuse)
    assert comm._p2p["ok"], comm._p2p.get("error")
    xs = rng.integers(-1000, 1000, 4099 * world).astype(np.int64)
    X = dist.ShardedVector.from_global(comm, xs)
    for it in range(50):
        assert int(X.sum()) == int(xs.sum()) and int(dist.dot(X, X)) == int((xs * xs).sum())
    fs = rng.uniform(-1, 1, 100003).astype(np.float32)
    F = dist.ShardedVector.from_global(comm, fs)
    first = F.sum()
    parts = [fs[slice(*dist.partition(fs.size, world, r))] for r in range(world)]
    assert all(F.sum() == first for _ in range(5))                      # deterministic
    assert abs(float(first) - float(fs.astype(np.float64).sum())) <= 2.0 ** -20 * float(np.abs(fs).sum())
    comm.barrier()
    td.destroy_process_group()
    print("rank %d/%d ok" % (rank, world))


if __name__ == "__main__":
    main()
