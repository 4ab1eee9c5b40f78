import sys, os
sys.path.insert(0, "/root/repo")
import torch
from totalsegmentator2d_b200 import _lib as L, synth
from totalsegmentator2d_b200.core.inference.predictor import B200nnUNetPredictor
def run(nst):
    feats=[32,64,128,256,512,512,512,512][:nst]
    arch=synth.make_arch_kwargs(feats)
    plans=synth.make_plans(arch,(512,512),(1.5,1.5),["ZScoreNormalization"]*2)
    dj=synth.make_dataset_json(["mean","max"],[f"l{i}" for i in range(26)])
    sd=synth.init_state_dict(arch,2,26,1003)
    p=B200nnUNetPredictor(precision="fp16",max_batch=12)
    p.manual_initialization(plans,"2d",dj,[sd],(0,1))
    x=torch.randn(12,512,512,2,device="cuda").half()
    st=L.stream_ptr()
    for _ in range(3): L.check(p._lib.ts2d_unet_forward(p._handles[0],x.data_ptr(),12,st))
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): L.check(p._lib.ts2d_unet_forward(p._handles[0],x.data_ptr(),12,st))
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/20
    print(nst,"stages: forward %.3f ms, launches per forward %d"%(ms, p.launch_count()//23))
    p.close()
for n in (8,7,6,5): run(n)
