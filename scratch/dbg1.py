import sys; sys.path.insert(0,".")
import numpy as np, torch, oracle
sys.path.insert(0,'.')
from ipex_llm_b200.transformers import low_bit_linear as lbl
rng=np.random.default_rng(0)
w=(rng.standard_normal((96,512))*0.02).astype(np.float32)
got=lbl.ggml_convert_qtype(torch.from_numpy(w),2,device="cpu").numpy().reshape(-1,34)
want=oracle.quantize(w,"sym_int4").reshape(-1,34)
bad=np.argwhere(got!=want)
print("mismatch bytes",len(bad),"blocks",len(set(bad[:,0])))
from collections import Counter
print(Counter(bad[:,1]).most_common(10))
for b,c in bad[:8]:
    blk=w.reshape(-1,64)[b]
    print(b,c,got[b,c],want[b,c], "d bytes", got[b,:2], want[b,:2])
    if c>=2:
        j=c-2
        d=np.frombuffer(want[b,:2].tobytes(),dtype=np.float16)[0]
        amax_i=np.abs(blk).argmax(); dd=np.float32(blk[amax_i])/np.float32(-8); idd=np.float32(1)/dd
        print("  x0",blk[j],"x1",blk[j+32],"x0*id+8.5",np.float32(blk[j]*idd)+np.float32(8.5),np.float32(blk[j+32]*idd)+np.float32(8.5))
