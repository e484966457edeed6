"""Renders small frames with the library RTC_BUILD_DIR selects and prints counters + a hash (A/B of item scheduling)."""
import hashlib, importlib, json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent.parent))
rtc = importlib.import_module("ray-tracing-c_b200")
for scene_id, width, spp in [(0, 64, 8), (0, 200, 100), (1, 160, 40), (6, 100, 33), (7, 80, 16)]:
    sess = rtc.Session(scene_id, width=width, spp=spp, depth=50)
    scene = rtc.Scene(sess.flatten())
    rgb8, rad, st = scene.render(sess.camera, rtc.make_opts(kernel="mega", rng="sample"))
    print(json.dumps({"scene": scene_id, "w": width, "spp": spp, "samples": st["samples"], "want": sess.width * sess.height * spp,
                      "segments": st["segments"], "sha": hashlib.sha256(rad.tobytes()).hexdigest()[:12]}), flush=True)
    scene.close()
