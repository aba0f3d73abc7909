import subprocess, sys, tempfile, os
sys.path.insert(0, "/root/repo")
import torch, numpy as np
from gru_mpi_cuda_b200 import GRUConfig
from gru_mpi_cuda_b200.models import oracle
from gru_mpi_cuda_b200.utils import weights as W
cfg = GRUConfig(); p = W.random_init(cfg, 5)
d = tempfile.mkdtemp(); wf = os.path.join(d, "w.bin"); of = os.path.join(d, "o.bin")
W.save(wf, p)
N, seed = 128, 77
r = subprocess.run(["baseline/_ref/namegen_ref", wf, str(N), str(seed), "1", "0", of], capture_output=True, text=True, env=dict(os.environ, RANK="0", WORLD_SIZE="1"))
print(r.returncode, r.stdout[-300:], r.stderr[-300:])
ref = torch.from_numpy(np.fromfile(of, dtype=np.uint8).reshape(N, cfg.max_len + 1).copy())
want = oracle.generate(cfg, p, W.random_floats(N * cfg.max_len, seed))
print("equal rows", (ref == want).all(dim=1).float().mean().item())
print("first char equal", (ref[:, 0] == want[:, 0]).float().mean().item(), "ref<=oracle", (ref[:, 0] <= want[:, 0]).float().mean().item())
print("mean first ref", ref[:, 0].float().mean().item(), "oracle", want[:, 0].float().mean().item())
print(ref[:4].tolist()); print(want[:4].tolist())
