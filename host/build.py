"""Builds the plain-C host side: host/libchannel_host.so (conf + driver, used by the tests through
ctypes) and host/channel_b200 (the run.conf-driven executable), both linked against
channel_gpu_b200/lib/libchannel_b200.so."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIBDIR = os.path.join(HERE, "..", "channel_gpu_b200", "lib")


def build_host():
    srcs = [os.path.join(HERE, f) for f in ("conf.c", "driver.c", "h5min.c", "checkpoint.c")]
    common = ["-O2", "-std=c99", "-Wall", "-D_POSIX_C_SOURCE=200809L", "-fPIC"]
    link = ["-L" + LIBDIR, "-lchannel_b200", "-Wl,-rpath," + os.path.abspath(LIBDIR), "-lm"]
    so = os.path.join(HERE, "libchannel_host.so")
    exe = os.path.join(HERE, "channel_b200")
    subprocess.run(["gcc"] + common + ["-shared", "-o", so] + srcs + link, check=True)
    subprocess.run(["gcc"] + common + ["-o", exe, os.path.join(HERE, "main.c")] + srcs + link, check=True)
    return so, exe


if __name__ == "__main__":
    print(build_host())
