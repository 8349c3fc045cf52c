import importlib, json, sys
sys.path.insert(0,'/root/repo')
import torch, bench
b = importlib.import_module("3dreconstruction_b200")
eng = b.Engine()
print(json.dumps(bench.ransac_extras(torch, b, eng, bench.peaks()), indent=1))
