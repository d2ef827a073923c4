"""Mirror of the record-level part of pkg/collector/ringbuf.go over the GPU path: a dump of
40-byte llm_slo_event records becomes packed lsg_event records on the device, ready for ProcessBatch."""
from dataclasses import dataclass

from . import _abi as A
from ._lib import default_context

RECORD_BYTES = 40
# signalFromType, ringbuf.go:199-226
SIGNAL_TYPES = {1: ("dns_latency_ms", "ms"), 2: ("tcp_retransmits_total", "count"), 3: ("runqueue_delay_ms", "ms"),
                4: ("connect_latency_ms", "ms"), 5: ("tls_handshake_ms", "ms"), 6: ("cpu_steal_pct", "ns"),
                7: ("mem_reclaim_latency_ms", "ms"), 8: ("disk_io_latency_ms", "ms"), 9: ("syscall_latency_ms", "ms")}
STATUS = ["ok", "warning", "error"]   # signals/generator.go:234-242


def signalFromType(st: int):
    return SIGNAL_TYPES.get(st, ("unknown", "unknown"))


@dataclass
class DecodedBatch:
    events: object     # numpy array of lsg_event
    status: object     # uint8 codes into STATUS
    errno: object      # int32


def DecodeBatch(raw: bytes, ident: int = 0, trace: int = 0, ts_offset_ns: int = 0, ctx=None) -> DecodedBatch:
    """decodeBPFEvent + toProbeEvent for every record of `raw` (ringbuf.go:152-197)"""
    ctx = ctx or default_context()
    ev, st, er = ctx.decode_ringbuf(raw, ident, trace, ts_offset_ns)
    return DecodedBatch(ev, st, er)


def signal_name(trace_sig: int) -> str:
    s = trace_sig & 15
    return A.SIGNALS[s] if s < A.NSIG else "unknown"
