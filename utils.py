"""Metric schema, validators and start-up logging helpers (reference utils.py:15-63)."""
from __future__ import annotations

import logging
import sys
from typing import Dict, List, Tuple

from pydantic import BaseModel, StrictFloat, confloat, conint

from dalle_b200.dht import BytesWithPublicKey, RecordValidatorBase, RSASignatureValidator, SchemaValidator
from dalle_b200.transport.p2p import Multiaddr, choose_ip_address


def get_logger(name=None) -> logging.Logger:
    return logging.getLogger(name)


def use_log_handler(level=logging.INFO):
    """Colourised single-line records on the root logger (the reference installs hivemind's handler)."""
    root = logging.getLogger()
    if not any(getattr(h, "_dalle_b200", False) for h in root.handlers):
        h = logging.StreamHandler(sys.stderr)
        h._dalle_b200 = True
        h.setFormatter(logging.Formatter("\033[2m%(asctime)s\033[0m [%(levelname)s] [%(name)s] %(message)s", "%b %d %H:%M:%S"))
        root.addHandler(h)
    root.setLevel(level)


logger = get_logger(__name__)


class LocalMetrics(BaseModel):
    """One peer's contribution record under DHT key ``{prefix}_metrics`` (strictly typed)."""

    step: conint(ge=0, strict=True)
    samples_per_second: confloat(ge=0.0, strict=True)
    samples_accumulated: conint(ge=0, strict=True)
    loss: StrictFloat
    mini_steps: conint(ge=0, strict=True)

    # pydantic-1 spellings used by the reference (callback.py:83, run_aux_peer.py:110)
    def dict(self, **kw):  # noqa: A003
        return self.model_dump(**kw)

    @classmethod
    def parse_obj(cls, obj):
        return cls.model_validate(obj)


class MetricSchema(BaseModel):
    metrics: Dict[BytesWithPublicKey, LocalMetrics]


def make_validators(experiment_prefix: str) -> Tuple[List[RecordValidatorBase], bytes]:
    signer = RSASignatureValidator()
    return [SchemaValidator(MetricSchema, prefix=experiment_prefix), signer], signer.local_public_key


BOLD, BLUE, RESET = "\033[1m", "\033[34m", "\033[0m"


def log_visible_maddrs(visible_maddrs: List[Multiaddr], only_p2p: bool) -> None:
    """Print the ``--initial_peers ...`` line other peers should use to join."""
    maddrs = [m if isinstance(m, Multiaddr) else Multiaddr(m) for m in visible_maddrs]
    if only_p2p:
        line = " ".join(f"/p2p/{pid}" for pid in dict.fromkeys(m.peer_id for m in maddrs))
    else:
        with_ip = [m for m in maddrs if m.host]
        if with_ip:
            best = choose_ip_address(with_ip)
            chosen = [m for m in maddrs if m.host == best]
        else:
            chosen = maddrs
        line = " ".join(str(m) for m in chosen)
    logger.info(f"Running a DHT peer. To connect other peers to this one, use {BOLD}{BLUE}--initial_peers {line}{RESET}")
    logger.info(f"Full list of visible multiaddresses: {' '.join(str(m) for m in maddrs)}")


def log_process_rank(training_args) -> None:
    logger.info(f"Process rank: {training_args.local_rank}, device: {training_args.device}, "
                f"n_gpu: {training_args.n_gpu}, distributed training: {training_args.local_rank != -1}, "
                f"16-bits training: {training_args.fp16}")
