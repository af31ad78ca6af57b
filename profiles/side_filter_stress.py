import importlib, sys, json, numpy as np, torch
sys.path.insert(0,'.')
import bench
cvr = importlib.import_module("master-thesis_b200")
dev=torch.device("cuda",0); torch.cuda.set_device(0)
r = bench._side_filter_stress(cvr, torch, dev, 0, int(sys.argv[1]) if len(sys.argv)>1 else 20_000_000, 6552.6)
print(json.dumps({k:v for k,v in r.items() if k!='clocks'}))
