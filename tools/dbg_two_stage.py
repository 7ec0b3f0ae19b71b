import importlib,sys,numpy as np
sys.path.insert(0,'.')
rtb=importlib.import_module('rust-ray-tracing_b200')
S,M=rtb.Sphere,rtb.Material
base=rtb.scenes.random_spheres()
W,H=192,108
cam=rtb.Camera.default(W,H)
r=rtb.CudaRenderer(device=0,seed=1,rays_per_lane=2)
def diff(scene,name):
    out={}
    for st in (1,2):
        r.stages=st
        img,lin,stat=r.render(1,1,scene,cam,want_linear=True)
        out[st]=lin.copy()
    d=np.abs(out[1][...,:3]-out[2][...,:3]).max(-1)
    print(name,len(scene),"diff pixels",int((d>1e-6).sum()))
mats={"m":M.lambertian((0.5,0.5,0.5))}
big=S((-4.0,1.0,0.0),1.0,"m"); ground=S((0.0,-1000.0,0.0),1000.0,"m")
small=[S((float(i%8)-4,0.2,float(i//8)-4),0.2,"m") for i in range(64)]
diff(rtb.Scene.from_list([ground,big],mats),"ground+big")
diff(rtb.Scene.from_list([big],mats),"big only")
diff(rtb.Scene.from_list([ground]+small[:31]+[big],mats),"33 spheres big last")
diff(rtb.Scene.from_list([ground]+small[:40]+[big],mats),"42 spheres big last")
diff(rtb.Scene.from_list([big,ground]+small[:40],mats),"42 spheres big first")
objs=list(base.objects)
diff(rtb.Scene(base.materials,objs[-3:]+objs[:-3]),"random, big first")
diff(rtb.Scene(base.materials,objs[:-3]),"random, no big")
diff(rtb.Scene(base.materials,objs[:256]+objs[-3:]),"random 259")
print("---- big sphere at index k of 486")
small_objs=objs[1:-3]
for k in (31,32,200,230,255,256,300,448,479,480,481,482,483,484,485):
    o2=[objs[0]]+small_objs[:k-1]+[objs[-2]]+small_objs[k-1:]+[small_objs[0],small_objs[1]]
    o2=o2[:486]
    diff(rtb.Scene(base.materials,o2),f"k={k}")
for n in (480,481,482,484,485,486,487,500,512):
    o2=([objs[0]]+small_objs+small_objs)[:n-1]+[objs[-2]]
    diff(rtb.Scene(base.materials,o2),f"big last of n={n}")
