import importlib, sys, json, torch
sys.path.insert(0,'.')
import bench
cvr = importlib.import_module("master-thesis_b200")
torch.cuda.set_device(0)
r = bench._side_png(cvr, torch, torch.device("cuda",0), 0)
print(json.dumps({k:v for k,v in r.items() if k!='clocks'}))
