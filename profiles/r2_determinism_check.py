import sys; sys.path.insert(0, str(__import__("pathlib").Path(__file__).resolve().parents[1]))
import numpy as np, gsdf_loader, importlib
pkg=gsdf_loader.load(); synth=importlib.import_module("gradient_sdf_b200.synth")
W,H=640,480
frames,poses,K,_=synth.make_sequence(100,W,H)
nest=pkg.NormalEstimator(W,H,K,11)
for cls,vs,tr in ((pkg.MapGradPixelSdf,0.02,10),(pkg.MapPixelSdf,0.01,5)):
    sdf=cls(np.float32(vs),np.float32(tr)*np.float32(vs),nest,log2_capacity=17)
    for i in range(4): sdf.update(frames[i],poses[i])
    opt=pkg.RigidPointOptimizer(sdf)
    out=[]
    for rep in range(20):
        opt.set_pose(poses[3]); c=opt.optimize(frames[4]); out.append((c,opt.last_iterations,opt.pose().tobytes()))
    print(cls.__name__, "distinct results over 20 runs:", len(set(out)), out[0][:2])
    sdf.close()
