# drivers.train over configurations the committed tests do not visit: base_filters 16 / 64, batch 1, the perceptual term with the
# recorded step, fp32 mode, augmentation on uint8 batches, a ragged last batch at every one of them.  Prints the epoch losses; a
# run is "ok" when every loss is finite and the training loss of the last epoch is below the first one's.  Run through gpurun.
import io, json, sys, tempfile
import torch
sys.path.insert(0, '.')
from oracle import losses_oracle as LO
from oracle.synth import make_pairs
from mri_superresolution_b200.drivers import TrainConfig, train
dev = torch.device('cuda')
bad = 0
CASES = [
    dict(name="f16 batch 4 + ragged 3", f=16, bs=4, n=11, hw=(32, 48)),
    dict(name="f64 batch 2 + ragged 1", f=64, bs=2, n=5, hw=(64, 64)),
    dict(name="f32 batch 1", f=32, bs=1, n=4, hw=(40, 24)),
    dict(name="f32 perceptual 0.1, graph", f=32, bs=3, n=8, hw=(32, 32), pw=0.1),
    dict(name="f64 perceptual mse, graph", f=64, bs=2, n=4, hw=(48, 32), pw=0.2, ptype="mse"),
    dict(name="f32 fp32 mode", f=32, bs=2, n=5, hw=(24, 40), prec="fp32"),
    dict(name="f32 no graph", f=32, bs=4, n=9, hw=(32, 32), graph=False),
    dict(name="f32 uint8 batches + augmentation", f=32, bs=4, n=10, hw=(64, 64), u8=True),
]
for c in CASES:
    h, w = c["hw"]
    lr, hr = make_pairs(c["n"] + 4, h, w, seed=len(c["name"]))
    if c.get("u8"):
        lr, hr = (lr[:, 0] * 255).to(torch.uint8), (hr[:, 0] * 255).to(torch.uint8)
    tr = [(lr[i:i + c["bs"]], hr[i:i + c["bs"]]) for i in range(0, c["n"], c["bs"])]
    va = [(lr[c["n"]:], hr[c["n"]:])]
    with tempfile.TemporaryDirectory() as d:
        cfg = TrainConfig(epochs=4, learning_rate=2e-3, ssim_weight=0.4, perceptual_weight=c.get("pw", 0.0),
                          perceptual_loss_type=c.get("ptype", "l1"), base_filters=c["f"], patience=10, checkpoint_dir=d, seed=1,
                          use_graph=c.get("graph", True), precision=c.get("prec", "bf16"),
                          vgg_weights=LO.make_vgg_params(seed=1235) if c.get("pw") else None, batch_size=c["bs"],
                          augmentation=bool(c.get("u8")))
        buf = io.StringIO()
        try:
            out = train(cfg, lambda: tr, lambda: va, device=dev, log_stream=buf)
            tl = [e["train_loss"] for e in out["history"]]
            vl = [e["val_loss"] for e in out["history"]]
            ok = all(x == x and abs(x) < 1e6 for x in tl + vl) and tl[-1] < tl[0]
            bad += not ok
            print(f"{c['name']:36s} train {' '.join('%.4f' % x for x in tl)} | val {' '.join('%.4f' % x for x in vl)} {'ok' if ok else 'PROBLEM'}", flush=True)
        except Exception as ex:
            bad += 1
            print(f"{c['name']:36s} FAILED {type(ex).__name__}: {str(ex)[:300]}", flush=True)
print("problems:", bad)
