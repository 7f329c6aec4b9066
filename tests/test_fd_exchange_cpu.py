"""CPU tier: the file-descriptor exchange the VMM symmetric heap uses to hand cuMemCreate allocations between the
rank processes (parallel/symm.py FdExchange: SCM_RIGHTS over abstract unix sockets), run with 2 and 3 gloo ranks.
Each rank shares a pipe's write end; every peer writes its rank into it; the owner reads W-1 distinct bytes."""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_SCRIPT = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, %r)
    import torch.distributed as dist
    from pretraining_llm_b200.parallel.symm import FdExchange
    dist.init_process_group("gloo")
    r, w = dist.get_rank(), dist.get_world_size()
    ex = FdExchange.get(r, w)
    rd, wr = os.pipe()
    fds = ex.all_to_all(wr)
    for q, fd in enumerate(fds):
        if q != r:
            os.write(fd, bytes([r]))
            os.close(fd)
    dist.barrier()
    got = sorted(os.read(rd, 64))
    assert got == [q for q in range(w) if q != r], (r, got)
    # second round on the same sockets: root -> everybody
    rd2, wr2 = os.pipe()
    f = ex.from_root(wr2 if r == 0 else None)
    if r != 0:
        os.write(f, bytes([100 + r]))
    dist.barrier()
    if r == 0:
        assert sorted(os.read(rd2, 64)) == [100 + q for q in range(1, w)]
    dist.barrier()
    if r == 0:
        print("FD_EXCHANGE_OK")
    dist.destroy_process_group()
""") % ROOT


@pytest.mark.parametrize("world", [2, 3])
def test_fd_exchange_between_processes(world, tmp_path):
    script = tmp_path / "fdx.py"
    script.write_text(_SCRIPT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(29810 + world), str(script)],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "FD_EXCHANGE_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]
