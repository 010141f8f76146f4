"""Node encoder time at small row counts (development aid): python tools/enc_small.py [rows ...]"""
import importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
lib_mod = importlib.import_module("gnn-cca_b200._lib")
lib = lib_mod.load()
dev = torch.device("cuda:0")
l1, l2 = torch.nn.Linear(2048, 128).to(dev), torch.nn.Linear(128, 32).to(dev)
for rows in [int(v) for v in sys.argv[1:]] or [2304, 2560, 5000, 10000]:
    xs = [torch.relu(torch.randn(rows, 2048, device=dev)) for _ in range(8)]
    h = torch.empty(rows, 32, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    def run(i):
        lib_mod.check(lib.gnncca_node_encoder(xs[i % 8].data_ptr(), l1.weight.data_ptr(), l1.bias.data_ptr(), l2.weight.data_ptr(), l2.bias.data_ptr(),
                                              h.data_ptr(), rows, 2048, 128, 32, lib_mod.MATH_TF32, st), "enc")
    for i in range(10): run(i)
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for i in range(200): run(i)
    z.record(); torch.cuda.synchronize()
    print("rows=%d min_rows=%s: %.2f us" % (rows, os.environ.get("GNNCCA_ENC_MIN_ROWS", "64"), a.elapsed_time(z) / 200 * 1e3))
