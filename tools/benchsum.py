import signal; signal.signal(signal.SIGPIPE, signal.SIG_DFL)
import sys,json
d=json.loads([l for l in sys.stdin.read().strip().splitlines() if l.startswith('{')][-1])
print("value %.1f ms/step %.2f e2e %.1f launches %d" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["gpu_launches"]))
print(" serialized", d.get("serialized_pass"), {k:(round(v["ms"]/2,2),v["launches"]//2) for k,v in d["kernel_ms"].items()})
print(" roofline", {k:d["roofline"][k] for k in ("kernel","achieved","frac","avg_launch_ms","mrays_per_s_in_kernel")}, d["traversal_work"])
