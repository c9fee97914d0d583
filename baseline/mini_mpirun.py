#!/usr/bin/env python
"""mini_mpirun: starts N ranks of a shim-MPI program on this host (harness for the reference).

    python baseline/mini_mpirun.py -np N [--gpu-per-rank] [--shm-bytes B] prog [args...]

Every rank gets B2I_MPI_RANK / B2I_MPI_SIZE / B2I_MPI_SHM; with --gpu-per-rank rank r also gets
CUDA_VISIBLE_DEVICES=r, which spreads the reference's `cudaSetDevice(0)` (kernel.cu:147) over the
GPUs without touching its source. Without it all ranks share GPU 0, as the reference is written.
"""
import argparse
import os
import subprocess
import sys


def launch(np_, prog, args=(), gpu_per_rank=False, shm_bytes=256 << 20, env=None, capture=True, timeout=600):
    name = f"/b2i_mpi_{os.getpid()}_{int.from_bytes(os.urandom(3), 'little')}"
    path = "/dev/shm" + name
    with open(path, "wb") as f:
        f.truncate(64 + shm_bytes)
    procs = []
    try:
        for r in range(np_):
            e = dict(os.environ if env is None else env)
            e.update(B2I_MPI_RANK=str(r), B2I_MPI_SIZE=str(np_), B2I_MPI_SHM=name, B2I_MPI_SHM_BYTES=str(shm_bytes))
            if gpu_per_rank:
                e["CUDA_VISIBLE_DEVICES"] = str(r)
            procs.append(subprocess.Popen([prog, *args], env=e, stdout=subprocess.PIPE if capture else None,
                                          stderr=subprocess.PIPE if capture else None, text=True))
        outs = []
        for p in procs:
            try:
                o, er = p.communicate(timeout=timeout)
            except subprocess.TimeoutExpired:
                p.kill()
                o, er = p.communicate()
            outs.append((p.returncode, o or "", er or ""))
        return outs
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
        try:
            os.unlink(path)
        except OSError:
            pass


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("-np", type=int, default=1)
    ap.add_argument("--gpu-per-rank", action="store_true")
    ap.add_argument("--shm-bytes", type=int, default=256 << 20)
    ap.add_argument("prog")
    ap.add_argument("args", nargs="*")
    a = ap.parse_args()
    res = launch(a.np, a.prog, a.args, a.gpu_per_rank, a.shm_bytes, capture=False)
    sys.exit(max(rc or 0 for rc, _, _ in res))
