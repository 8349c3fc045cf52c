import importlib, json, sys
sys.path.insert(0,'/root/repo')
import torch, bench
b = importlib.import_module("3dreconstruction_b200")
eng = b.Engine()
print(json.dumps(bench.plugin_call_extras(torch, b, eng), indent=1))
