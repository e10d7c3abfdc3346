import os, sys, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import test_mega as t
m = t._tiny(128)
from torchao_fork_b200._models.llama import Generator
try:
    gen = Generator(m, 1, 64, temperature=0.0, top_k=None, use_graph=False, mega=True)
    gen.engine.set_state(torch.tensor([3], device="cuda"), 0)
    gen.engine.step(); torch.cuda.synchronize()
    print("status", gen.engine.mega_status(), gen.engine.logits[0,:8])
except Exception as e:
    print("ERR", e)
