#!/usr/bin/env python
"""Development probe: the end-to-end arm's copy pattern alone (no kernels): per picture 14 uploads of its batch sizes on
one stream, the picture download as three plane copies on another; each direction alone and both at once."""
import torch


def main(up_sizes=None, label="the bench picture's 15 uploads"):
    up_sizes = up_sizes or [4925472, 64 * 240, 8192, 2600000, 2552424, 2527588, 2048, 8192, 3 * 518400, 3 * 518400, 3 * 518400, 3060,
                20400, 6120, 9000]
    dn_sizes = [3840 * 2160 * 2, 1920 * 1080 * 2, 1920 * 1080 * 2]
    h_up = [torch.empty(n, dtype=torch.uint8).pin_memory() for n in up_sizes]
    d_up = [torch.empty(n, dtype=torch.uint8, device="cuda") for n in up_sizes]
    h_dn = [torch.empty(n, dtype=torch.uint8).pin_memory() for n in dn_sizes]
    d_dn = [torch.empty(n, dtype=torch.uint8, device="cuda") for n in dn_sizes]
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()

    def run(up, dn, reps=60):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s_up.wait_event(e0)
        s_dn.wait_event(e0)
        for _ in range(reps):
            if up:
                with torch.cuda.stream(s_up):
                    for d, h in zip(d_up, h_up):
                        d.copy_(h, non_blocking=True)
            if dn:
                with torch.cuda.stream(s_dn):
                    for d, h in zip(d_dn, h_dn):
                        h.copy_(d, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s_up)
        torch.cuda.current_stream().wait_stream(s_dn)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    run(True, True, 5)
    print("==", label)
    print("bytes per picture: up %.1f MB, down %.1f MB" % (sum(up_sizes) / 1e6, sum(dn_sizes) / 1e6))
    print("uploads alone   %.3f ms per picture" % run(True, False))
    print("download alone  %.3f ms per picture" % run(False, True))
    print("both            %.3f ms per picture" % run(True, True))


if __name__ == "__main__":
    main()
    main([17300000], "one upload")
    main([17300000 // 2] * 2, "two uploads")
    main([17300000 // 6] * 6, "six equal uploads")
    main([17000000] + [20000] * 14, "one big + 14 small")
    main([4925472, 2600000, 2552424, 2527588, 3 * 518400, 3 * 518400, 3 * 518400], "only the seven big ones")
