import json,sys
d=json.loads(sys.stdin.read())
print(d["stress_1080p_forward"]["fps"], {k:v["us"] for k,v in d["stress_1080p_forward"]["stages"].items()}, d["sort_2p24"]["ms"])
