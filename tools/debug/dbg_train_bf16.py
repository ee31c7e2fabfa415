import sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import xvector_b200
from oracle import numpy_ref, synth, train_ref
cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG)
dim, speakers, lr, step = 24, 101, 0.05, 100
batch, frames = int(sys.argv[1]), int(sys.argv[2])
w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=77), speakers, bias=False)
x, y = synth.config5_batch(batch, frames, dim, speakers, seed=78)
p = numpy_ref.OracleParams(**cfg)
mom = float(cfg["batchnorm_momentum"])
new_b, info_b = train_ref.train_step(x, y, w, p, lr, step, bf16_gemm=True)
new_e, info_e = train_ref.train_step(x, y, w, p, lr, step)
for prec in ("bf16",):
    eng = xvector_b200.Engine(cfg, dim, precision=prec); eng.set_weights(w); eng.train_init(speakers)
    loss, total = eng.train_step(x, y, lr, step)
    print(prec, "loss gpu %.7f oracle_bf16 %.7f exact %.7f" % (loss, info_b["loss"], info_e["loss"]))
    for i in range(1, 8):
        for st in ("moving_mean", "moving_variance"):
            name = "tdnn/tdnn%d_bn/%s" % (i, st)
            v = eng.train_weight(name, np.asarray(w[name]).shape)
            w0 = np.asarray(w[name], np.float64)
            mu_g = (v - w0 * mom) / (1 - mom)
            mu_b = (new_b[name] - w0 * mom) / (1 - mom)
            mu_e = (new_e[name] - w0 * mom) / (1 - mom)
            print("  %s: |gpu-bf16o| %.3e  |gpu-exact| %.3e  |bf16o-exact| %.3e  scale %.3e" % (
                name, np.abs(mu_g - mu_b).max(), np.abs(mu_g - mu_e).max(), np.abs(mu_b - mu_e).max(), np.abs(mu_e).max()))
    for name in sorted(info_b["grads"]):
        g = eng.train_grad(name, np.asarray(w[name]).shape)
        gb = info_b["grads"][name].reshape(g.shape); ge = info_e["grads"][name].reshape(g.shape)
        print("  grad %-28s vs bf16o %.3e  vs exact %.3e  (bf16o vs exact %.3e)" % (
            name, np.linalg.norm(g - gb) / np.linalg.norm(gb), np.linalg.norm(g - ge) / np.linalg.norm(ge), np.linalg.norm(gb - ge) / np.linalg.norm(ge)))
    eng.close()
