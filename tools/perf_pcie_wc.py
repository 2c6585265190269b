#!/usr/bin/env python
"""Development probe: does write-combined pinned memory on the upload side change what both directions get at once?"""
import ctypes as C
import torch


def main():
    n = 24883200
    rt = C.CDLL("libcudart.so")
    torch.zeros(1, device="cuda")
    d_up = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_dn = torch.empty(n, dtype=torch.uint8, device="cuda")
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()

    def host(flags):
        p = C.c_void_p()
        assert rt.cudaHostAlloc(C.byref(p), C.c_size_t(n), C.c_uint(flags)) == 0
        return p

    for label, fl in (("default", 0), ("write-combined upload", 4), ("portable", 1)):
        h_up, h_dn = host(fl), host(0)
        C.memset(h_up, 1, n)

        def run(up, dn, reps=40):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            s_up.wait_event(e0)
            s_dn.wait_event(e0)
            for _ in range(reps):
                if up:
                    rt.cudaMemcpyAsync(C.c_void_p(d_up.data_ptr()), h_up, C.c_size_t(n), 1, C.c_void_p(s_up.cuda_stream))
                if dn:
                    rt.cudaMemcpyAsync(h_dn, C.c_void_p(d_dn.data_ptr()), C.c_size_t(n), 2, C.c_void_p(s_dn.cuda_stream))
            torch.cuda.current_stream().wait_stream(s_up)
            torch.cuda.current_stream().wait_stream(s_dn)
            e1.record()
            torch.cuda.synchronize()
            return n * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9
        run(True, True, 5)
        print("%-24s H2D alone %.1f  D2H alone %.1f  both %.1f GB/s each" % (label, run(True, False), run(False, True), run(True, True)))


if __name__ == "__main__":
    main()
