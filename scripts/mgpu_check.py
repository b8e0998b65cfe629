"""Multi-GPU check under torchrun: the load-balanced driver over NCCL must reproduce, bit for bit, what each rank
gets when it integrates all of its own cells locally (the reference's LB == non-LB criterion,
tests/src/standardChemistryModel-Test.C:90-102, here across real GPUs with real redistribution)."""
import os, sys
from pathlib import Path
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from loadbalancedchemistrymodel_b200 import capi, workloads as W
from loadbalancedchemistrymodel_b200.model import CudaCellSolver, LoadBalancedChemistryModel

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
mech, Y, T, p, rho = W.gri_skewed(n + 17 * rank, rank=rank, world=world, hot_fraction=0.4 / world)
chem = capi.Chemistry(mech, capi.default_options(solver="seulex", abs_tol=1e-12, rel_tol=1e-1), device=local)
tY, tT, tp, trho = (torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (Y, T, p, rho))
nloc = tT.shape[0]
lb = LoadBalancedChemistryModel(mech.n_species, CudaCellSolver(chem), dev)
ref = LoadBalancedChemistryModel(mech.n_species, CudaCellSolver(chem), dev); ref.world = 1
dtc_lb = dtc_ref = torch.full((nloc,), 1e-7, dtype=torch.float64, device=dev)
ok = True
for step in range(3):
    m1 = lb.solve(tY, tT, tp, trho, 1e-6, dtc_lb); m2 = ref.solve(tY, tT, tp, trho, 1e-6, dtc_ref)
    same = (torch.equal(lb.RR, ref.RR) and torch.equal(lb.deltaTChem, ref.deltaTChem) and torch.equal(lb.T_out, ref.T_out)
            and torch.equal(lb.cell_cost, ref.cell_cost) and torch.equal(lb.status, ref.status) and m1 == m2)
    ok &= bool(same)
    print(f"rank {rank} step {step}: identical={bool(same)} total_cost={lb.total_cost:.0f} sent={lb.last['sent_cells']} recv={lb.last['recv_cells']} plan={lb.send_plan}", flush=True)
    dtc_lb, dtc_ref = lb.deltaTChem, ref.deltaTChem
flag = torch.tensor([1 if ok else 0], device=dev); dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0: print("MGPU_CHECK", "PASS" if int(flag.item()) == 1 else "FAIL", flush=True)
chem.close(); dist.destroy_process_group()
