#!/usr/bin/env python
"""Mutation fuzz of the OBJ / MTL importer and the scene-file reader (host/model.cpp, host/serializer.cpp) under ASan + UBSan.  CPU only.
    cd path_tracer_b200/host && g++ -std=c++17 -O1 -g -fsanitize=address,undefined -pthread -I . -I ../../include -o /tmp/asan_host \
        ../../tools/fuzz_scene_loaders_main.cpp scene.cpp pack.cpp spectrum.cpp procedural.cpp model.cpp image.cpp serializer.cpp -lz
    python tools/fuzz_scene_loaders.py
Round 2: 2000 + 2100 (polygon-rich seeds) mutated OBJ files and 1500 mutated scene files, no sanitizer report (one finding fixed: enumerators read from a
scene file are range-checked now)."""
import sys, os, numpy as np, tempfile, subprocess
sys.path.insert(0,"/root/repo"); sys.path.insert(0,"/root/repo/tests")
from path_tracer_b200 import scene as S, _build
import test_reference_host as T
S.ensure_spectrum_table()
rng=np.random.default_rng(3)
d=tempfile.mkdtemp()
T.write_obj(os.path.join(d,"m.obj"), True)
obj=open(os.path.join(d,"m.obj"),"rb").read()
# a saved scene
sc=S.Scene.make("cornell"); sc.save(os.path.join(d,"s.json")); js=open(os.path.join(d,"s.json"),"rb").read()
print("seed sizes", len(obj), len(js), os.listdir(d))
env=dict(os.environ, PTB_SPECTRUM_TABLE=_build.SPECTRUM_TABLE)
def run(kind, seed, ext, n, textual):
    files=[]
    for it in range(n):
        s=bytearray(seed)
        for _ in range(rng.integers(1,8)):
            m=rng.integers(0,4)
            if m==0 and len(s)>0:
                s[rng.integers(0,len(s))]= rng.choice(list(b"0123456789-. /\n e")) if textual and rng.random()<0.7 else rng.integers(0,256)
            elif m==1 and len(s)>16: s=s[:rng.integers(8,len(s))]
            elif m==2:
                i=rng.integers(0,max(1,len(s))); s[i:i]=bytes(rng.choice(list(b"0123456789-/ fv\n"),size=rng.integers(1,10)).tolist())
            else:
                i=rng.integers(0,max(1,len(s)-20)); j=i+rng.integers(1,20); del s[i:j]
        p=os.path.join(d,"f_%s_%05d%s"%(kind,it,ext)); open(p,"wb").write(bytes(s)); files.append(p)
    bad=0
    for k in range(0,len(files),200):
        r=subprocess.run(["/tmp/asan_host",kind]+files[k:k+200],stdout=subprocess.PIPE,stderr=subprocess.PIPE,text=True,cwd=d,timeout=600)
        if r.returncode!=0 or "runtime error" in r.stderr or "AddressSanitizer" in r.stderr:
            bad+=1; print(kind,"BAD batch",k, r.stderr[:1800]); break
        print(kind, r.stdout.strip())
    return bad
b1=run("obj",obj,".obj",2000,True)
# quads and larger polygons (the triangulation of host/model.cpp), relative indices, several shapes and materials
sys.path.insert(0,"/root/repo/tools")
import fuzz_obj
for k in range(3):
    fuzz_obj.write_random_obj(os.path.join(d,"p%d.obj"%k), np.random.default_rng(40+k))
    b1+=run("obj",open(os.path.join(d,"p%d.obj"%k),"rb").read(),".obj",700,True)
b2=run("scene",js,".json",1500,True)
print("bad",b1,b2)
