import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d["roofline"]["kernels"]; print(d["ms_per_step"], {n:k[n]["ms"] for n in k if n.startswith("step")})
