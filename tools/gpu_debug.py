import sys, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'tests'); sys.path.insert(0,'tools')
import nats_llm_studio_b200 as b200, orc, make_gguf
np.set_printoptions(linewidth=200)
e = b200.Engine(max_ctx=1024)
g = np.load('tests/golden/tiny-llama.npz'); p='tests/golden/tiny-llama.gguf'
toks = g['tokens']
m = e.load_file('dbg', p)
ref = orc.Model(p, act_quant=1, kv_f16=1).forward(toks)
for trial in range(3):
    got = m.eval(toks)
    d = np.abs(got-ref).max(axis=1)/np.abs(ref).max()
    print("trial", trial, np.round(d,5))
t2 = np.arange(1, 70, dtype=np.int32)
ref2 = orc.Model(p, act_quant=1, kv_f16=1).forward(t2)
got2 = m.eval(t2)
print(np.round(np.abs(got2-ref2).max(axis=1)/np.abs(ref2).max(),5))
