import sys, os
sys.path.insert(0, "svt-av1-psyex_b200/python"); sys.path.insert(0, "tests")
from checkers import Oracle
from helpers import PictureSet
from me_b200.engine import Engine
ps = PictureSet(320, 192, 17, "motion", Oracle())
eng = Engine(320, 192, 4, 3, pictures_in_flight=24, devices=[0, 1])
print("init ok", flush=True)
for k in (0, 1, 2):
    eng.register_picture(k, ps.get(k).planes[0], 320, 192)
    print("registered", k, flush=True)
eng.close()
