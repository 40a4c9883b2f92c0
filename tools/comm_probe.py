"""Multi-GPU plumbing probe: run under torchrun (one rank per GPU).  Exercises mf_comm_create / mf_comm_selftest:
VMM allocation + POSIX-fd exchange, peer mappings, the device flag barrier and NVLink read/write bandwidth.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/comm_probe.py
"""
import ctypes as C
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mantaflow_b200 import _lib  # noqa: E402


def main():
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    session = "probe%s" % os.environ.get("MASTER_PORT", "0")
    l = _lib.lib()
    c = l.mf_comm_create(rank, world, local, session.encode())
    if not c:
        raise SystemExit("rank %d: mf_comm_create failed: %s" % (rank, l.mf_last_error().decode()))
    out = (C.c_double * 16)()
    mb = int(os.environ.get("PROBE_MB", "256"))
    _lib.check(l.mf_comm_selftest(c, mb << 20, 2000, out))
    print(json.dumps({"rank": rank, "world": world, "mismatches": out[0], "barrier_us": out[1], "peer_read_gbs": out[2],
                      "peer_write_gbs": out[3], "granularity": out[4], "local_copy_gbs": out[5], "slab_mb": mb,
                      "chain_us_plain": out[6], "chain_us_own_slab": out[7], "chain_us_peer_slab": out[8]}), flush=True)
    l.mf_comm_destroy(c)


if __name__ == "__main__":
    main()
