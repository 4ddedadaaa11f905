import sys; sys.path.insert(0,'/root/repo')
import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi
with vr.Renderer(64,64) as r:
    r.generate_volume(size=512, format=capi.VOL_R16F, layout=capi.LAYOUT_TEXTURE)
    print("tex_stream GB/s", r.measure_tex_stream())
