"""Phase times of the fused peer exchange kernel (build csrc/optim.cu with -DNMB_PEER_CLOCKS): run under torchrun with
2+ ranks; prints per rank the mean ns of the phases of CTA 100 and of the whole kernel."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402


def main():
    world = int(os.environ["WORLD_SIZE"])
    rank = int(os.environ["RANK"])
    sys.argv = [sys.argv[0], "--gpus", str(world), "--steps", "400", "--warmup", "5", "--no-cpu-baseline", "--no-selfcheck",
                "--sustained-s", "0"]
    orig_finish = bench._finish

    def finish(w):
        from nonmyopia_b200 import _lib
        out = (C.c_uint64 * 12)()
        _lib.lib().nmb_debug_peer_clocks(out)
        n, k = max(out[3], 1), max(out[10], 1)
        print(f"rank {rank}: CTA 100 over {out[3]} launches: reduce+push {out[0] / n:.0f} ns, bias corrections + receive "
              f"{out[1] / n:.0f} ns, sum+AdamW {out[2] / n:.0f} ns | whole kernel {out[7] / k:.0f} ns, first CTA start -> "
              f"last CTA's AdamW {out[8] / k:.0f} ns", file=sys.stderr, flush=True)
        orig_finish(w)

    bench._finish = finish
    bench.main()


if __name__ == "__main__":
    main()
