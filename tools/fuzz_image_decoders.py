#!/usr/bin/env python
"""Mutation fuzz of host/image.cpp (PNG / JPEG / Radiance HDR / TGA / BMP / PNM / GIF / PSD / PIC decoders) under AddressSanitizer + UBSan.  CPU only.
    cd path_tracer_b200/host && g++ -std=c++17 -O1 -g -fsanitize=address,undefined -I . -I ../../include -o /tmp/asan_img ../../tools/fuzz_image_decoders_main.cpp image.cpp -lz
    python tools/fuzz_image_decoders.py
Round 2: 12 000 + 8 000 (TGA / BMP seeds) mutated files, no sanitizer report (two findings on the way, fixed: signed overflow in the IDCT on absurd
coefficients, a scan that names a Huffman table the file never defined)."""
import sys, os, numpy as np, tempfile, subprocess
sys.path.insert(0,"/root/repo"); sys.path.insert(0,"/root/repo/tests")
from PIL import Image
import test_image_decode as T
rng=np.random.default_rng(7)
d=tempfile.mkdtemp()
seeds=[]
a=T.picture(40,30,1)
for name,kw in (("a.png",{}),("b.jpg",dict(quality=80)),("c.jpg",dict(quality=60,progressive=True,subsampling=2)),("g.jpg",dict(quality=70,subsampling=1))):
    p=os.path.join(d,name); Image.fromarray(a[...,:3]).save(p,**kw); seeds.append(open(p,"rb").read())
Image.fromarray(a[...,:3]).quantize(16).save(os.path.join(d,"p.png")); seeds.append(open(os.path.join(d,"p.png"),"rb").read())
Image.fromarray((a[...,0].astype(np.uint16)*257)).save(os.path.join(d,"s.png")); seeds.append(open(os.path.join(d,"s.png"),"rb").read())
rgbe=rng.integers(0,256,size=(9,40,4)).astype(np.uint8); T.write_hdr(os.path.join(d,"h.hdr"),rgbe,True); seeds.append(open(os.path.join(d,"h.hdr"),"rb").read())
b=T.picture(40,30,1)
def add(name, im, **kw):
    p=os.path.join(d,name); im.save(p,**kw); seeds.append(open(p,"rb").read())
add("a.tga", Image.fromarray(b[...,:3])); add("b.tga", Image.fromarray(b), compression="tga_rle")
add("c.tga", Image.fromarray(b[...,:3]).quantize(17), compression="tga_rle"); add("d.tga", Image.fromarray(b[...,[0,3]],"LA"))
T.write_psd(os.path.join(d,"p.psd"), np.ascontiguousarray(b.transpose(2,0,1)), rle=True); seeds.append(open(os.path.join(d,"p.psd"),"rb").read()); T.write_psd(os.path.join(d,"q.psd"), np.ascontiguousarray(b.transpose(2,0,1)).astype(np.uint16)*257, depth=16); seeds.append(open(os.path.join(d,"q.psd"),"rb").read()); T.write_pic(os.path.join(d,"r.pic"), b, [(2,0xE0),(1,0x10)], rng); seeds.append(open(os.path.join(d,"r.pic"),"rb").read()); add("k.jpg", Image.fromarray(b,"CMYK"), quality=70); add("a.gif", Image.fromarray(b[...,:3]).quantize(40)); add("b.gif", Image.fromarray(b[...,:3]).quantize(7), interlace=True, transparency=3); seeds.append(b"P6\n40 30\n255\n"+b[...,:3].tobytes()); seeds.append(b"P5 40 30 65535\n"+b[...,:2].tobytes()); add("a.bmp", Image.fromarray(b[...,:3])); add("b.bmp", Image.fromarray(b)); add("c.bmp", Image.fromarray(b[...,:3]).quantize(17)); add("d.bmp", Image.fromarray(b[...,0]>128))
files=[]
for it in range(16000):
    s=bytearray(seeds[it%len(seeds)])
    for _ in range(rng.integers(1,6)):
        m=rng.integers(0,3)
        if m==0 and len(s)>0: s[rng.integers(0,len(s))]=rng.integers(0,256)
        elif m==1 and len(s)>8: s=s[:rng.integers(4,len(s))]
        else:
            i=rng.integers(0,max(1,len(s))); s[i:i]=bytes(rng.integers(0,256,size=rng.integers(1,8)).tolist())
    p=os.path.join(d,"f%05d.bin"%it); open(p,"wb").write(bytes(s)); files.append(p)
bad=0
for k in range(0,len(files),400):
    r=subprocess.run(["/tmp/asan_img"]+files[k:k+400],stdout=subprocess.PIPE,stderr=subprocess.PIPE,text=True)
    if r.returncode!=0 or "runtime error" in r.stderr or "AddressSanitizer" in r.stderr:
        bad+=1; print(r.stderr[:1500])
        break
    print(r.stdout.strip())
print("bad batches",bad)
