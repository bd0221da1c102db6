"""What storage does the box have besides tmpfs?  Mounts, block devices, and buffered / O_DIRECT write+read rates."""
import os, sys, time, mmap, subprocess, json
from concurrent.futures import ThreadPoolExecutor
for cmd in ("df -h -x proc -x sysfs -x devtmpfs 2>/dev/null | head -30", "lsblk 2>/dev/null | head -30", "grep -E ' / | /tmp | /dev/shm | /root' /proc/mounts | cut -c1-300"):
    print("$", cmd); print(subprocess.run(cmd, shell=True, capture_output=True, text=True).stdout)
N = 256 << 20; FILES = 8
buf = mmap.mmap(-1, N); buf.write(os.urandom(1 << 20) * (N >> 20))
def run(root, direct, write):
    flags = (os.O_WRONLY | os.O_CREAT | os.O_TRUNC) if write else os.O_RDONLY
    if direct: flags |= os.O_DIRECT
    def one(i):
        fd = os.open(os.path.join(root, f"probe{i}"), flags, 0o644)
        try:
            off = 0
            mv = memoryview(buf)
            while off < N:
                n = os.pwrite(fd, mv[off:off + (16 << 20)], off) if write else os.preadv(fd, [mv[off:off + (16 << 20)]], off)
                if n <= 0: break
                off += n
            if write: os.fsync(fd)
        finally: os.close(fd)
    t0 = time.monotonic()
    with ThreadPoolExecutor(FILES) as ex: list(ex.map(one, range(FILES)))
    return round(FILES * N / (time.monotonic() - t0) / 1e9, 2)
for root in sys.argv[1:]:
    os.makedirs(root, exist_ok=True); rec = {"root": root}
    for direct in (False, True):
        try:
            rec[f"write_gbps_direct{int(direct)}"] = run(root, direct, True)
            rec[f"read_gbps_direct{int(direct)}"] = run(root, direct, False)
        except OSError as e:
            rec[f"direct{int(direct)}_error"] = str(e)
    for i in range(FILES):
        try: os.remove(os.path.join(root, f"probe{i}"))
        except OSError: pass
    print(json.dumps(rec), flush=True)
