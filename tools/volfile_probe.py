"""Volume-file path timings on the GPU box: bake -> save (+pyramid) -> staged upload from the file (page cache hot),
next to vr_upload_volume from host memory and the device-side pyramid alone."""
import os, sys, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import volumetricrenderer_b200 as vr
from volumetricrenderer_b200 import capi, volfile

sizes = [int(a) for a in sys.argv[1:]] or [512, 1024]
tmp = tempfile.mkdtemp(prefix="volrt_")
with vr.Renderer(64, 64) as r:
    for n in sizes:
        gb = n ** 3 * 2 / 1e9
        r.generate_volume(size=n, format=capi.VOL_R16F, threshold=0.07)
        path = os.path.join(tmp, "smoke%d.vol" % n)
        t0 = time.perf_counter(); pyr = r.volume_pyramid(); t_pyr = time.perf_counter() - t0
        t0 = time.perf_counter(); r.save_volume_file(path, with_pyramid=True); t_save = time.perf_counter() - t0
        host = r.download_volume()
        res = {}
        for name, flags in (("plain", 0), ("verify", capi.FILE_VERIFY), ("trust pyramid", capi.FILE_TRUST_PYRAMID)):
            ts = []
            for _ in range(3):
                t0 = time.perf_counter(); r.upload_volume_file(path, capi.LAYOUT_TEXTURE, flags); ts.append(time.perf_counter() - t0)
            res[name] = min(ts)
        ts = []
        for _ in range(3):
            t0 = time.perf_counter(); r.upload_volume(host, capi.VOL_R16F, capi.LAYOUT_TEXTURE); ts.append(time.perf_counter() - t0)
        t_mem = min(ts)
        print("%d^3 fp16 (%.2f GB): pyramid %d levels %.1f ms incl. D2H | save+pyramid %.2f s | upload from file %.1f ms (%.1f GB/s), "
              "with SHA-256 %.1f ms, trusting the file's pyramid %.1f ms | vr_upload_volume from pageable host memory %.1f ms (%.1f GB/s)" %
              (n, gb, len(pyr), t_pyr * 1e3, t_save, res["plain"] * 1e3, gb / res["plain"], res["verify"] * 1e3, res["trust pyramid"] * 1e3,
               t_mem * 1e3, gb / t_mem), flush=True)
        os.remove(path)
os.rmdir(tmp)
