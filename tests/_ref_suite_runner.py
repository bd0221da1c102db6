"""Runs test files of the REFERENCE (read where they lie under /root/reference/tests, never copied) against
THIS build.  A throw-away directory gets a one-file package named `torchsnapshot` whose search path is this
repository's package first and the reference's package second, so `torchsnapshot.batcher`, `.io_preparer`,
`.scheduler`, `.snapshot`, ... are this build's modules, and only modules that are not part of the hot path
(test_utils, stateful, state_dict, ...) come from the reference.  The directory goes on PYTHONPATH so that
the worker processes the reference's multi-rank tests spawn resolve the same package.

    python tests/_ref_suite_runner.py <pytest args> /root/reference/tests/test_batcher.py

Used by tests/test_reference_suite.py in a subprocess, so the aliasing never leaks into the other tests."""
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("TSB200_REFERENCE_ROOT", "/root/reference")
PKG = os.path.join(ROOT, "torchsnapshot-1_b200")

_PACKAGE_INIT = '''import os as _os
__path__ = [{pkg!r}, {ref!r}]
_init = _os.path.join({pkg!r}, "__init__.py")
exec(compile(open(_init).read(), _init, "exec"))
'''

# pytest-asyncio is not installed here: run `async def` tests on a private event loop, and register the
# markers the reference's tests use
_ASYNC_SHIM = '''import asyncio
import inspect


def pytest_configure(config):
    for name in ("asyncio", "cpu_and_gpu", "gpu_only", "s3_integration_test", "gcs_integration_test"):
        config.addinivalue_line("markers", name + ": marker used by the reference's tests")


def pytest_sessionfinish(session, exitstatus):
    # which files the `torchsnapshot.*` modules were loaded from: the caller checks that the hot path is
    # this build's code, not the reference's
    import json
    import os
    import sys

    out = os.environ.get("TSB200_REFSUITE_MODULES")
    if out:
        mods = {{name: getattr(m, "__file__", None) for name, m in sys.modules.items()
                if name == "torchsnapshot" or name.startswith("torchsnapshot.")}}
        with open(out, "w") as f:
            json.dump(mods, f, indent=1, sort_keys=True)


def pytest_pyfunc_call(pyfuncitem):
    if not inspect.iscoroutinefunction(pyfuncitem.obj):
        return None
    kwargs = {{name: pyfuncitem.funcargs[name] for name in pyfuncitem._fixtureinfo.argnames}}
    loop = asyncio.new_event_loop()
    try:
        loop.run_until_complete(pyfuncitem.obj(**kwargs))
    finally:
        loop.close()
    return True
'''


def make_shim(directory: str) -> None:
    os.makedirs(os.path.join(directory, "torchsnapshot"), exist_ok=True)
    with open(os.path.join(directory, "torchsnapshot", "__init__.py"), "w") as f:
        f.write(_PACKAGE_INIT.format(pkg=PKG, ref=os.path.join(REF, "torchsnapshot")))
    with open(os.path.join(directory, "tsb_ref_async_shim.py"), "w") as f:
        f.write(_ASYNC_SHIM.format())


def run(pytest_args, timeout=None, modules_json=None) -> subprocess.CompletedProcess:
    with tempfile.TemporaryDirectory(prefix="tsb_refsuite_") as shim:
        make_shim(shim)
        env = dict(os.environ)
        if modules_json:
            env["TSB200_REFSUITE_MODULES"] = modules_json
        env["PYTHONPATH"] = shim + os.pathsep + env.get("PYTHONPATH", "")
        env["PYTHONDONTWRITEBYTECODE"] = "1"  # /root/reference is read-only
        cmd = [sys.executable, "-m", "pytest", "-p", "tsb_ref_async_shim", "-p", "no:cacheprovider", "-q",
               "--rootdir", shim, "-c", os.devnull] + list(pytest_args)
        return subprocess.run(cmd, env=env, cwd=shim, capture_output=True, text=True, timeout=timeout)


if __name__ == "__main__":
    res = run(sys.argv[1:])
    sys.stdout.write(res.stdout)
    sys.stderr.write(res.stderr)
    sys.exit(res.returncode)
