import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
E = importlib.import_module("ollama-operator_b200").engine
ctx = int(sys.argv[1]) if len(sys.argv) > 1 else 512
print(E.attn_bench(ctx, 32, 32, 128, 4, 32, 24) * 1e3, "us")
