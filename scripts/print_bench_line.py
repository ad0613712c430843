import json,sys
tag=sys.argv[1]
lines=[l for l in sys.stdin.read().splitlines() if l.startswith('{')]
d=json.loads(lines[-1])
print(tag, round(d["value"],1), round(d["ms_per_step"],4), round(d["roofline"]["avg_launch_ms"],4), d["gpu_launches"], round(d["e2e"]["value"],1), d["clocks"])
