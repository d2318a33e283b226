# Runs several rank scripts back to back inside ONE N-way launch (importing torch costs ~6 s per
# process on this image, so one launch per script would dominate the test time). Each script
# does its own Init ... Finalize, exactly like the reference's per-file jobs
# (/root/reference/test/runtests.jl:11-16).
import os, runpy, sys, traceback

here = os.path.dirname(os.path.abspath(__file__))
rank = os.environ.get("RANK", "0")
for name in sys.argv[1:]:
    argv = name.split(":")
    sys.argv = [os.path.join(here, argv[0])] + argv[1:]
    try:
        runpy.run_path(sys.argv[0], run_name="__main__")
    except SystemExit as e:
        if e.code not in (0, None):
            print(f"SCRIPT_FAIL {argv[0]} rank {rank}: exit {e.code}", flush=True)
            sys.exit(1)
    except BaseException:
        traceback.print_exc()
        print(f"SCRIPT_FAIL {argv[0]} rank {rank}", flush=True)
        sys.exit(1)
    print(f"SCRIPT_OK {argv[0]} rank {rank}", flush=True)
