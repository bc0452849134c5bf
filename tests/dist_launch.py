"""Spawn W local gloo ranks the way an MPI launcher would (env vars only)."""
import json
import os
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
def next_port():
    """A TCP port that is free right now (bind to 0 and let the kernel pick)."""
    import socket
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


_RENDEZVOUS_ERRORS = ("Address already in use", "Connection refused", "Connection reset", "timed out", "Timed out",
                      "failed to connect", "Broken pipe")


def launch(world, argv, script=None, extra_env=None, timeout=300, cwd=None):
    """Run `world` ranks; one retry on a fresh port if the TCP rendezvous itself failed (the free port found by
    next_port() can be taken by another process before rank 0 binds it)."""
    out = _launch_once(world, argv, script, extra_env, timeout, cwd)
    if any(out[1]) and any(e in log for log in out[2] for e in _RENDEZVOUS_ERRORS):
        out = _launch_once(world, argv, script, extra_env, timeout, cwd)
    return out


def _launch_once(world, argv, script=None, extra_env=None, timeout=300, cwd=None):
    out_dir = tempfile.mkdtemp(prefix="mdt_test_")
    port = next_port()
    procs = []
    for r in range(world):
        env = dict(os.environ)
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "SLURM_NPROCS", "SLURM_PROCID", "SLURM_NODELIST", "LSB_HOSTS"):
            env.pop(k, None)
        env.update(OMPI_COMM_WORLD_SIZE=str(world), OMPI_COMM_WORLD_RANK=str(r), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), DDP_BACKEND="gloo", OMP_NUM_THREADS="1", MDT_DROPIN="0",
                   PYTHONPATH=os.path.dirname(HERE) + os.pathsep + env.get("PYTHONPATH", ""))
        env.update(extra_env or {})
        cmd = [sys.executable, script or os.path.join(HERE, "_dist_worker.py")] + [str(a).replace("{out}", out_dir) for a in argv]
        procs.append(subprocess.Popen(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=cwd))
    outs = []
    for p in procs:
        try:
            o, _ = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        outs.append(o)
    codes = [p.returncode for p in procs]
    return out_dir, codes, outs


def results(out_dir, world):
    return [json.load(open(os.path.join(out_dir, f"rank{r}.json"))) for r in range(world)]
