"""Per-parameter gradient error of the training step vs the golden reference run (debug aid)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import stofnet_b200 as sb
for name in sys.argv[1:] or ["different-armadillo", "clean-serenity"]:
    g = dict(np.load(os.path.join(ROOT, "tests/golden/train_%s.npz" % name)))
    sd = {k: torch.from_numpy(v) for k, v in np.load(os.path.join(ROOT, "tests/golden/weights_%s.npz" % name)).items()}
    dev = torch.device("cuda:0")
    model = sb.build_model(sd, dev).train()
    x = torch.from_numpy(g["x"]).unsqueeze(1).to(dev)
    gt = torch.from_numpy(g["gt"]).to(dev)
    loss = sb.mask_loss(model(x), gt, 7, 1.0, 20.0, float(g["lambda"]))
    loss.backward()
    print(name, "loss", float(loss), g["losses"][0])
    for k, p in model.named_parameters():
        ref = g["grad0/" + k]; got = p.grad.cpu().numpy()
        print("  %-45s rel err %.3e   max|ref| %.3e" % (k, np.abs(got - ref).max() / (np.abs(ref).max() + 1e-30), np.abs(ref).max()))
