import os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "p-ntfields_b200"))
from pntf import synth
from pntf.model import Model_multi
from pntf.trainer import run_training
dev = torch.device("cuda:0")
d, rows = 3, 60000
datas, Bs = [], []
for e in range(2):
    c, s = synth.sample_pairs(rows, d, dev, seed=60 + e, margin=0.5 / 12)
    datas.append(torch.cat([c, s], dim=1)); Bs.append(synth.sample_B(d, dev, seed=70 + e, E=1)[0])
torch.manual_seed(0)
model = Model_multi("", "", d, 2, device=str(dev)); model.init_network()
for call in range(2):
    h = run_training(model, datas, Bs, epochs=4, log=print)
    print("call", call, h, float(model.network.flat().flat.abs().sum()), model.optimizer.step_count)
