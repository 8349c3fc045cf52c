import importlib, json, sys
sys.path.insert(0, '/root/repo')
import torch, bench
b = importlib.import_module("3dreconstruction_b200")
eng = b.Engine()
r = bench.essential_extras(torch, b, eng, bench.peaks()[0])["ransac_essential_1024_pairs"]
r.pop("note", None)
print(json.dumps(r))
