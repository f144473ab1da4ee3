import json,sys
d=json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
print(sys.argv[1], "ms/frame %.4f e2e %.4f serial %.4f call %.4f"%(d["ms_per_frame"], d["e2e"]["ms_per_frame"], d["latency"]["device_ms_per_frame_single_stream"], d["latency"]["match_call_ms_host_to_host"]))
print({k:round(v*1000,2) for k,v in d["roofline"]["kernel_ms_warm"].items()})
