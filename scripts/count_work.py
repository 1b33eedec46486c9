import importlib, sys
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rt = importlib.import_module("ray-tracer_b200")
dev = rt.Device(0)
for name,w,h,spp,d in [("book-one",960,540,8,50),("cornell-box",600,600,8,50),("book-two",400,400,8,50),("stress:1000000",960,540,4,16)]:
    sc = rt.Scene.example(dev,name,1)
    cam = sc.flat.camera(w/h)
    s,_,st = sc.render(cam,w,h,spp,d,flags=rt.RENDER_COUNT_WORK, want_sumsq=False)
    s,_,st2 = sc.render(cam,w,h,spp,d,flags=rt.RENDER_STAGE_TIMING, want_sumsq=False)
    print(name, sc.bvh_info(), "rays/path %.2f box/ray %.1f prim/ray %.1f"%(st.rays/st.paths, st.box_tests/st.rays, st.prim_tests/st.rays),
          "ms %.2f (g %.2f e %.2f s %.2f a %.2f) iters %d Grays/s %.2f in-extend %.2f"%(st2.ms_total, st2.ms_generate, st2.ms_extend, st2.ms_shade, st2.ms_accumulate, st2.iterations, st2.rays/st2.ms_total/1e6, st2.rays/st2.ms_extend/1e6))
