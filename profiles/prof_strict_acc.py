"""Strict-mode accuracy against the fp64 run of the reference (tests/golden/score64_*.npz) per precision mode and per
accumulation-group length: python profiles/prof_strict_acc.py   (B200_TC32_FLUSH is read at import: one process per value)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import cases
from b_pinn_kalman_filter_b200.models import utils as mutils
dev = torch.device("cuda", 0)
for name in ("cifar10_ncsnpp_ve", "nc_ddpm_vp"):
    cfg, sd, x, labels, out = cases.score_case(name)
    out64 = cases.score_fp64(name)
    row = ["%s: reference fp32 vs fp64 %.2e" % (name, cases.rel_l2(out, out64))]
    for precision in ("fp32_simt", "fp32", "tf32"):
        m = mutils.ScoreModel(cfg, device=dev, precision=precision, init=False)
        m.load_state_dict(sd)
        y = m(x.to(dev), labels.to(dev)).cpu()
        row.append("%s %.2e (max %.2e)" % (precision, cases.rel_l2(y, out64), cases.max_rel(y, out64)))
    print("flush=%s  " % os.environ.get("B200_TC32_FLUSH", "default") + "; ".join(row))
