import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
from conftest import golden_cases, load_golden
from test_gpu_train import run_train_case
from test_gpu_mpn import close
dev = torch.device("cuda:0")
for case in [c for c in golden_cases("mpn_") if c.endswith("_train")]:
    z = load_golden("mpn_" + case)
    model, out, loss = run_train_case(z, dev)
    print(case, "loss", loss.item(), float(z["loss"]))
    for name, p in model.named_parameters():
        ref = torch.from_numpy(z["grad/" + name])
        print("   %-50s err %.2e  |ref|max %.2e" % (name, close(p.grad, ref), ref.abs().max()))
    name = "MPNet.node_model.node_mlp.fc_layers.0.weight"
    g, ref = model.state_dict()[name], torch.from_numpy(z["grad/" + name])
    g = dict(model.named_parameters())[name].grad.cpu()
    K = ref.shape[1] - 6
    print("   node block err", (g[:, :K] - ref[:, :K]).abs().max().item(), "edge block err", (g[:, K:] - ref[:, K:]).abs().max().item())
    print("   got edge block row0", g[0, K:].tolist()); print("   ref edge block row0", ref[0, K:].tolist())
    print("   got edge block col0", g[:4, K].tolist()); print("   ref col0", ref[:4, K].tolist())
