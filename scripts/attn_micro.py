"""Stand-alone timing of the decode-attention kernel through b2_attention_decode_static (CUDA events, warm and
cold L2): context lengths x split counts, llama3-8b head geometry."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "llama-fastapi-k8s-gpu_b200"))
import torch
from llama_cpp import _cabi
lib = _cabi.load_library()
n_head, n_kv, hd, bt = 32, 8, 128, 64
dev = torch.device("cuda:0")
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
stream = torch.cuda.current_stream().cuda_stream
for ctx in (150, 500, 2000):
    nblk = (ctx + bt - 1) // bt + 1
    k = torch.randn(nblk, n_kv, bt, hd, device=dev).half()
    v = torch.randn(nblk, n_kv, bt, hd, device=dev).half()
    q = torch.randn(n_head, hd, device=dev)
    out = torch.empty(n_head, hd, device=dev)
    table = torch.arange(nblk, dtype=torch.int32, device=dev)
    pos = torch.tensor([ctx - 1], dtype=torch.int32, device=dev)
    for ns in (2, 4, 8):
        def call():
            rc = lib.b2_attention_decode_static(0, q.data_ptr(), k.data_ptr(), v.data_ptr(), table.data_ptr(), nblk, pos.data_ptr(),
                                                n_head, n_kv, hd, bt, ns, out.data_ptr(), stream)
            assert rc == 0, rc
        for _ in range(5): call()
        res = {}
        for mode in ("warm", "cold"):
            ts = []
            for _ in range(20):
                if mode == "cold": flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); call(); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1) * 1e3)
            ts.sort(); res[mode] = ts[len(ts) // 2]
        print(f"ctx {ctx:5d} splits {ns}: warm {res['warm']:7.2f} us  cold {res['cold']:7.2f} us", flush=True)
