import sys; sys.path.insert(0,'.')
import volumerenderer_b200 as vr
r = vr.VolumeRenderer(0)
for k in (0,3,5,6,7,8,9,1): print(k, r.ubench(k))
