"""Per-step error table of the giant-graph parity check (bench.giant_parity) with the worst elements located."""
import copy, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from oracle import mpn_oracle, mpn_oracle_chunked as moc
gnncca = importlib.import_module("gnn-cca_b200"); engine = importlib.import_module("gnn-cca_b200.engine"); cfgmod = importlib.import_module("gnn-cca_b200.config")
C_, D_ = int(sys.argv[1]), int(sys.argv[2])
math = sys.argv[3] if len(sys.argv) > 3 else "tf32"
dev = torch.device("cuda:0")
wl = dict(bench.WORKLOADS["cfg5"], cams=C_, dets=D_)
params = cfgmod.graph_net_params()
torch.manual_seed(0)
model = gnncca.MOTMPNet(copy.deepcopy(params), None, "resnet50", math=math).to(dev).eval()
N, deg = C_ * D_, C_ * D_ - D_
x, ei, ea = bench.giant_inputs(wl, (0, N), dev)
with torch.no_grad():
    outs, lat, _ = engine.forward_inference(model, x, ei, ea, return_latents=True)
spec = mpn_oracle.MPNSpec(params, "resnet50")
sd64 = moc.cast_state_dict({k: v.detach().cpu() for k, v in model.state_dict().items()}, torch.float64)
g = torch.Generator().manual_seed(7)
rows = torch.sort(torch.randperm(N, generator=g)[:32])[0]
sel = (rows[:, None] * deg + torch.arange(deg)[None, :]).reshape(-1); sd_ = sel.to(dev)
row, col = ei[0, sd_].cpu(), ei[1, sd_].cpu()
h0 = lat[0][0].cpu().double(); e0 = lat[0][1][sd_].cpu().double()
def rel(a, b): return ((a.double() - b).abs() / (1 + b.abs()))
with torch.no_grad():
    e_ref = mpn_oracle.mlp_forward(sd64, "encoder.edge_mlp", spec.enc_edge, ea[sd_].cpu().double())
    r = rel(lat[0][1][sd_].cpu(), e_ref); print("enc edge", float(r.max()))
    for t in range(1, 5):
        e_new, h_new, lg = moc.motmpnet_step(spec, sd64, lat[t-1][0].cpu().double(), h0, lat[t-1][1][sd_].cpu().double(), e0, row, col, rows, classify=t >= 2)
        re_ = rel(lat[t][1][sd_].cpu(), e_new); rh = rel(lat[t][0][rows.to(dev)].cpu(), h_new)
        i = int(re_.argmax()); j = int(rh.argmax())
        print("step", t, "e err", float(re_.max()), "at", i // 6, i % 6, "got", float(lat[t][1][sd_].cpu().view(-1)[i]), "ref", float(e_new.view(-1)[i]),
              "| h err", float(rh.max()), "got", float(lat[t][0][rows.to(dev)].cpu().view(-1)[j]), "ref", float(h_new.view(-1)[j]),
              "| |h| max", float(h_new.abs().max()), "|e| max", float(e_new.abs().max()))
        if lg is not None:
            print("     logits err", float(rel(outs[t-2][sd_].cpu(), lg).max()))
