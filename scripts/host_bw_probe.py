"""Host DRAM write / copy bandwidth with the library's own host threads, for several thread counts
("host_bw with 4, 8 and 32 threads on the 8-GPU box").  One fresh process per count: the pool is sized once."""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
counts = [int(x) for x in sys.argv[1:]] or [4, 8, 16, 32]
code = ("import json, sys; sys.path.insert(0, %r); from quemb_b200 import eri_sparse_DF as q; "
        "w, c = q.host_bandwidth(2 << 30); print(json.dumps({'write_gbs': w, 'copy_gbs': c}))" % str(ROOT))
out = {"cores": len(os.sched_getaffinity(0))}
for n in counts:
    env = dict(os.environ, QEMB_HOST_THREADS=str(n))
    env.pop("LOCAL_WORLD_SIZE", None)
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    out[f"threads_{n}"] = json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else {"error": r.stderr[-300:]}
print(json.dumps(out))
