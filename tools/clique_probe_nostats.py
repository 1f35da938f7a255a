import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "clique_mix_probe.py")).read()
src = src.replace("s.step(P); ctx.synchronize(); s.status()", "s.set_stats(False); s.step(P); ctx.synchronize(); s.status()")
exec(compile(src, "probe", "exec"))
