import sys; sys.path.insert(0,'/root/repo')
import numpy as np, torch, gpu_raytracer_b200 as g, oracle as O
for shape in (0, 256*16+2, 128*16+2, 128*16+4, 256*16+4, 64*16+4):
    g.set_param("build.walk", shape)
    for n,kb in [(2,30),(3,30),(255,30),(257,63),(1025,30),(40001,30),(200000,63)]:
        t=O.gen_scene(O.SCENE_SOUP,n,0,seed=n); ob=O.OracleBvh(t,kb)
        with g.Bvh.build(torch.from_numpy(t).cuda(),key_bits=kb,keep_keys=True) as b:
            l,r,p=b.read_topology(); bx=b.read_boxes()
            assert np.array_equal(l,ob.left) and np.array_equal(r,ob.right) and np.array_equal(p,ob.parent),(shape,n)
            assert np.array_equal(bx.view(np.uint32),ob.boxes.view(np.uint32)),(shape,n)
    dup=np.repeat(O.gen_scene(O.SCENE_SOUP,300,0,seed=2),40,axis=0); ob=O.OracleBvh(dup,30)
    with g.Bvh.build(torch.from_numpy(dup).cuda()) as b:
        l,r,p=b.read_topology(); assert np.array_equal(l,ob.left) and np.array_equal(p,ob.parent)
    print("shape",shape,"ok",flush=True)
