import sys, torch
sys.path.insert(0,'.')
from tests import util
import guava_renderer_b200 as grb
DEV='cuda:0'
for (P,H,W,V,sm) in ((300000,1024,1024,4,1.0),(100000,512,512,8,1.0)):
    views=[util.inputs_for(P,H,W,4000+v,view=v,num_views=4,scale_mult=sm) for v in range(V)]
    st=grb.BatchRasterizationSettings(H,W,torch.tensor([[x["tanfovx"],x["tanfovy"]] for x in views],device=DEV),torch.zeros(V,32,device=DEV),1.0,torch.stack([x["viewmatrix"] for x in views]).to(DEV),torch.stack([x["projmatrix"] for x in views]).to(DEV))
    stack=lambda k: torch.stack([x[k] for x in views]).to(DEV)
    args=(stack("xyz"),torch.zeros(V,P,3,device=DEV),stack("features"),stack("opacity"),stack("scaling"),stack("rotation"),st)
    with torch.no_grad():
        for _ in range(3): grb.rasterize_gaussians_batch(*args)
        torch.cuda.synchronize()
        grb.set_option("profile_stages",1); grb._lib_mod.stage_times_reset()
        for _ in range(5): out=grb.rasterize_gaussians_batch(*args)
        torch.cuda.synchronize(); grb.set_option("profile_stages",0)
    print(P,H,W,V,{k:round(v[0]/max(v[1],1),4) for k,v in grb._lib_mod.stage_times().items()})
