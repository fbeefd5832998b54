import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, ROOT + "/tests")
import torch
from common import rel_l2, score_inputs
from ditsep_b200.diffsep import DiffSepModel, default_config
from oracle import ncsnpp_ref as R
xt, t, mixn = score_inputs(1, 7810)
for scale in (1.0, 0.3, 0.1, 0.01):
    cfg = R.ScoreConfig()
    sd = R.make_state_dict(cfg.backbone, seed=0, init_scale=scale)
    with torch.no_grad():
        ref = R.score_forward_ref(sd, cfg, xt, t, mixn)
    for prec in ("fp32", "bf16"):
        m = DiffSepModel(default_config(precision=prec)); m.score_model.load_state_dict(sd); m = m.to("cuda").eval()
        with torch.no_grad():
            out = m(xt.cuda(), t.cuda(), mixn.cuda())
        print(f"init_scale={scale} {prec}: rel-L2 {rel_l2(out, ref):.3e}", flush=True)
