import sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import xvector_b200
from oracle import numpy_ref, synth, train_ref
cfg = dict(synth.TDNN_AMSOFTMAX_CONFIG, loss_func="additive_margin_softmax", last_layer_linear=False, last_layer_no_bn=False, network_relu_type=sys.argv[1] if len(sys.argv)>1 else "prelu")
if cfg["network_relu_type"]=="relu": del cfg["network_relu_type"]
dim, speakers, batch, frames, lr = 24, 101, 8, 50, 0.05
seed=int(sys.argv[2]) if len(sys.argv)>2 else 31
w = synth.add_softmax_weights(synth.make_tdnn_weights(dim, cfg, seed=seed), speakers)
x, y = synth.config5_batch(batch, frames, dim, speakers, seed=seed+1)
eng = xvector_b200.Engine(cfg, dim, precision="fp32"); eng.set_weights(w); eng.train_init(speakers)
loss, total = eng.train_step(x, y, lr, 30000)
new_ref, info = train_ref.train_step(x, y, w, numpy_ref.OracleParams(**cfg), lr, 30000)
print(loss, info["loss"])
for name in ["tdnn/tdnn6_bn/beta","tdnn/tdnn6_dense/kernel","tdnn/tdnn5_relu/alpha","tdnn/tdnn5_bn/gamma","tdnn/tdnn5_bn/beta","tdnn/tdnn5_dense/bias","tdnn/tdnn5_dense/kernel","tdnn/tdnn4_bn/gamma","tdnn/tdnn4_bn/beta","tdnn/tdnn1_bn/beta"]:
    if name not in info["grads"]: continue
    g_ref = info["grads"][name]; g = eng.train_grad(name, np.asarray(w[name]).shape).reshape(g_ref.shape)
    d = np.abs(g-g_ref)
    i = np.argmax(d)
    print(name, "norm_ref %.3e normerr %.3e maxerr %.3e at %d: %.6e vs %.6e; n_bad(>1e-4*max)=%d" % (np.linalg.norm(g_ref), np.linalg.norm(g-g_ref), d.max(), i, g.flat[i], g_ref.flat[i], (d>1e-4*np.abs(g_ref).max()).sum()))
