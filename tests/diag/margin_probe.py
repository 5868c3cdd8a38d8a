"""Margins of the SNR thresholds asserted by the train() drop-in tests: three repetitions, printed."""
import json, os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import inr_for_audio_b200 as pkg
root = ROOT
for rep in range(3):
    tmp=tempfile.mkdtemp()
    kw = dict(inst="gt_bach", duration=1, method="wave", num_hidden_features=64, num_sine=2, num_snake=0, num_tanh=0,
              num_freq=32, omega=30, hidden_omega=30, learning_rate=1e-3, data_dir=os.path.join(root, "data"), seed=0, verbose=False)
    ck1 = pkg.train(tmp, "rff", total_steps=150, **kw)
    p1 = json.load(open(os.path.join(os.path.dirname(ck1), "parameters.json")))
    kw.update(num_hidden_features=128)
    ck3 = pkg.train(tmp, "rff4", total_steps=100, precision="fp16", **kw)
    p3 = json.load(open(os.path.join(os.path.dirname(ck3), "parameters.json")))
    common = dict(inst="gt_bach", duration=1, method="wave", num_snake=0, num_tanh=0, omega=30, total_steps=60, data_dir=os.path.join(root, "data"), seed=0, verbose=False)
    ck = pkg.train(tmp, "mlp", num_sine=2, num_hidden_features=64, **common)
    p0 = json.load(open(os.path.join(os.path.dirname(ck), "parameters.json")))
    ckk = pkg.train(tmp, "kan", arch="kan", num_hidden_features=32, **common)
    pk = json.load(open(os.path.join(os.path.dirname(ckk), "parameters.json")))
    print("rff fp32 SNR %.2f | rff fp16 SNR %.2f | mlp60 SNR %.3f | kan60 SNR %.3f" % (p1["SNR"], p3["SNR"], p0["SNR"], pk["SNR"]), flush=True)
