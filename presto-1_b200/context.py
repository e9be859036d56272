"""GpuContext: one pg_ctx (device + stream).  Creation fails loudly without a CUDA device (no CPU fallback)."""
import ctypes as C

from . import _native as N


class GpuContext:
    def __init__(self, device=0, use_torch_stream=True):
        self._h = C.c_void_p()
        rc = N.lib().pg_ctx_create(device, C.byref(self._h))
        if rc != N.PG_OK:
            raise N.PrestoGpuError(rc, "pg_ctx_create(device=%d) failed: no usable CUDA device (there is no CPU fallback)" % device)
        self.device = device
        if use_torch_stream:
            try:
                import torch
                if torch.cuda.is_available():
                    torch.cuda.set_device(device)
                    self.set_stream(torch.cuda.current_stream(device).cuda_stream)
            except ImportError:
                pass

    @property
    def handle(self):
        return self._h

    def set_stream(self, cuda_stream):
        self.stream = int(cuda_stream)
        N.check(self._h, N.lib().pg_ctx_set_stream(self._h, C.c_void_p(cuda_stream)))

    def reserve_sms(self, sms):
        """Leave `sms` multiprocessors free for a second pipeline stage running on another stream."""
        N.check(self._h, N.lib().pg_ctx_reserve_sms(self._h, int(sms)))

    def synchronize(self):
        N.check(self._h, N.lib().pg_ctx_synchronize(self._h))

    def launch_count(self):
        return int(N.lib().pg_ctx_launch_count(self._h))

    def set_timing(self, enabled):
        N.check(self._h, N.lib().pg_ctx_set_timing(self._h, 1 if enabled else 0))

    def kernel_report(self, reset=True):
        """{kernel name: (total ms, launches)} recorded since the last reset (CUDA events on the context stream)."""
        buf = C.create_string_buffer(1 << 16)
        N.check(self._h, N.lib().pg_ctx_kernel_report(self._h, buf, len(buf)))
        out = {}
        for item in buf.value.decode().split(";"):
            if item:
                name, rest = item.split("=")
                ms, cnt = rest.split("/")
                out[name] = (float(ms), int(cnt))
        if reset:
            N.check(self._h, N.lib().pg_ctx_kernel_time(self._h, None, 1, None, None))
        return out

    def close(self):
        if self._h:
            N.lib().pg_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_default = {}


def default_context(device=0):
    if device not in _default:
        _default[device] = GpuContext(device)
    return _default[device]
