import sys, os, importlib, numpy as np, torch
sys.path.insert(0, os.getcwd())
from oracle import lifter_oracle as lo
from oracle.philox import dropout_keep_mask
pkg = importlib.import_module("3d_poseestimation_b200")
def rel(a,b):
    a,b=torch.as_tensor(a).double().reshape(-1),torch.as_tensor(b).double().reshape(-1)
    return float((a-b).norm()/(b.norm()+1e-30))
for tag in ["small","small_nobn","h1024","wide"]:
    g=dict(np.load(f"tests/golden/lifter_{tag}.npz"))
    i_dim,o_dim,H,S,B,BN,seed=[int(v) for v in g["cfg"]]; p=float(g["p"])
    params,buffers=lo.numpy_init(i_dim,o_dim,H,S,seed)
    masks=[dropout_keep_mask(B,H,s,0,seed+7,p) for s in range(2*S+1)]
    X,Y=torch.from_numpy(g["x"]),torch.from_numpy(g["y"])
    ref=lo.lifter_forward_backward(params,buffers,X,lambda o:((o-Y)**2).mean(),num_stage=S,training=True,use_bn=bool(BN),p_dropout=p,masks=masks,need_input_grad=True)
    for fmt in ["fp16","bf16"]:
        m=pkg.LinearModel(i_dim,o_dim,linear_size=H,num_stage=S,p_dropout=p,BN=bool(BN)); m.load_state_dict({**params,**buffers}); m=m.cuda().train()
        m.train_format=fmt; m.dropout_seed,m.dropout_step=seed+7,0
        x=X.cuda().requires_grad_(True)
        out=m(x); loss=torch.nn.functional.mse_loss(out,Y.cuda()); loss.backward()
        print(f"== {tag} {fmt}: out {rel(out.detach().cpu(),ref['out']):.2e} dx {rel(x.grad.cpu(),ref['dx']):.2e}")
        dxe=(x.grad.cpu()-ref['dx']); 
        print("   dx per-row rel:", " ".join(f"{float(dxe[r].norm()/ref['dx'][r].norm()):.3f}" for r in range(min(B,16))))
        for name,prm in m.named_parameters():
            gr=ref['grads'][name]
            if prm.grad is None: print(f"   {name:40s} None"); continue
            print(f"   {name:40s} rel {rel(prm.grad.cpu(),gr):.3e}  |ref| {float(gr.norm()):.3e}")
