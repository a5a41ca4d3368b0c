import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
print(sys.argv[1], "fps", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "kernel_ms", round(r["kernel_ms_per_launch"],4), "frac", round(r["frac"],3), d["clocks"]["sm_mhz"])
if "mesh" in d: print("  mesh", round(d["mesh"]["mesh_ms"],3), "pc", round(d["mesh"]["point_cloud_ms"],3), d["mesh"].get("vertices"), d["mesh"].get("triangles"), d["mesh"].get("checksum"))
