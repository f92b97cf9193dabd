import faulthandler, sys, runpy, time, os
faulthandler.enable()
faulthandler.dump_traceback_later(int(os.environ.get('WD_SECS', '120')), exit=True)
sys.argv = ['bench.py'] + sys.argv[1:]
t0 = time.time()
try:
    runpy.run_path('bench.py', run_name='__main__')
except SystemExit as e:
    print("bench exited", e.code, "after", round(time.time() - t0, 1), "s", file=sys.stderr)
faulthandler.cancel_dump_traceback_later()
print("interpreter shutting down", file=sys.stderr)
