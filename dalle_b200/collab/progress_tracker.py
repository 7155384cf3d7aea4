"""Swarm-wide progress tracking towards ``target_batch_size`` (hivemind.optim.progress_tracker).

Each peer publishes a signed ``LocalTrainingProgress`` under DHT key ``{prefix}_progress`` (sub-key = its public
key) and periodically fetches everybody's records to derive ``GlobalTrainingProgress``: the swarm epoch (max), the
samples accumulated towards the current epoch (extrapolated by each peer's samples/s and record age), peer counts
and an ETA for the next epoch.  ``ready_to_update_epoch`` turns true when the target is reached or when this peer
is behind the swarm.
"""
from __future__ import annotations

import contextlib
import dataclasses
import threading
import time
from typing import Optional

from pydantic import BaseModel, StrictBool, StrictFloat, confloat, conint

from ..dht import DHT, BytesWithPublicKey, RSASignatureValidator, SchemaValidator, get_dht_time


class LocalTrainingProgress(BaseModel):
    peer_id: bytes
    epoch: conint(ge=0, strict=True)
    samples_accumulated: conint(ge=0, strict=True)
    samples_per_second: confloat(ge=0.0, strict=True)
    time: StrictFloat
    client_mode: StrictBool


class TrainingProgressSchema(BaseModel):
    progress: dict[BytesWithPublicKey, Optional[LocalTrainingProgress]]


@dataclasses.dataclass
class GlobalTrainingProgress:
    epoch: int
    samples_accumulated: int
    target_batch_size: int
    num_peers: int
    num_clients: int
    eta_next_epoch: float
    next_fetch_time: float


_PROCESS_KEY_OWNER: dict = {}   # which DHT node of this process signs with the process-wide key


class PerformanceEMA:
    """Exponential moving average of samples/s with bias correction; ``pause`` excludes global-step time."""

    def __init__(self, alpha: float = 0.1, eps: float = 1e-20, paused: bool = False):
        self.alpha, self.eps, self.num_updates = alpha, eps, 0
        self.ema_seconds_per_sample, self.samples_per_second = 0.0, eps
        self.timestamp = time.perf_counter()
        self.paused = paused
        self.lock = threading.Lock()

    def update(self, task_size: float, interval: Optional[float] = None) -> float:
        assert task_size > 0
        if not self.paused:
            now = time.perf_counter()
            if interval is None:
                interval = max(0.0, now - self.timestamp)
            self.timestamp = now
            self.ema_seconds_per_sample = self.alpha * interval / task_size + (1 - self.alpha) * self.ema_seconds_per_sample
            self.num_updates += 1
            adjusted = self.ema_seconds_per_sample / (1 - (1 - self.alpha) ** self.num_updates)
            self.samples_per_second = 1 / max(adjusted, self.eps)
        return self.samples_per_second

    def reset_timer(self):
        self.timestamp = time.perf_counter()

    @contextlib.contextmanager
    def pause(self):
        self.paused, was = True, self.paused
        try:
            yield
        finally:
            self.paused = was
            self.reset_timer()


class ProgressTracker(threading.Thread):
    def __init__(self, dht: DHT, prefix: str, target_batch_size: int, *, client_mode: Optional[bool] = None,
                 min_refresh_period: float = 0.5, max_refresh_period: float = 10, default_refresh_period: float = 3,
                 expected_drift_peers: float = 3, expected_drift_rate: float = 0.2, performance_ema_alpha: float = 0.1,
                 metadata_expiration: float = 60.0, status_loglevel=None, private_key=None, daemon: bool = True,
                 start: bool = True):
        super().__init__(name=f"{prefix}_tracker", daemon=daemon)
        self.dht, self.prefix = dht, prefix
        self.training_progress_key = f"{prefix}_progress"
        self.target_batch_size = target_batch_size
        self.client_mode = dht.client_mode if client_mode is None else client_mode
        self.min_refresh_period, self.max_refresh_period = min_refresh_period, max_refresh_period
        self.default_refresh_period = default_refresh_period
        self.expected_drift_peers, self.expected_drift_rate = expected_drift_peers, expected_drift_rate
        self.metadata_expiration = metadata_expiration
        self.performance_ema = PerformanceEMA(alpha=performance_ema_alpha)
        if private_key is None:
            # One signing key per DHT node: the key of a signature validator the node already has (two validators
            # with different keys would each demand their own envelope), else a fresh one.  Not the process-wide key:
            # two peers inside one process (tests, notebooks) would publish their progress under the same
            # [owner:...] sub-key and erase each other's record.
            private_key = getattr(dht, "_record_signing_key", None)
            if private_key is None:
                from ..dht.crypto import RSAPrivateKey

                existing = [v for v in getattr(getattr(dht, "_validator", None), "_validators", [])
                            if isinstance(v, RSASignatureValidator)]
                if existing:
                    private_key = existing[0]._private_key
                elif _PROCESS_KEY_OWNER.setdefault("dht", id(dht)) == id(dht):
                    private_key = RSAPrivateKey.process_wide()    # the first node of the process (the usual case)
                else:
                    private_key = RSAPrivateKey()
                dht._record_signing_key = private_key
        sig = RSASignatureValidator(private_key)
        self._local_public_key = sig.local_public_key
        dht.add_validators([SchemaValidator(TrainingProgressSchema, prefix=prefix), sig])
        self.local_progress = self._local(0, 0)
        self.global_progress = GlobalTrainingProgress(0, 0, target_batch_size, 1, int(self.client_mode),
                                                      float("inf"), get_dht_time() + default_refresh_period)
        self.lock_global, self.lock_local = threading.RLock(), threading.RLock()
        self.should_report, self.fetched = threading.Event(), threading.Event()
        self.shutdown_triggered, self.shutdown_complete = threading.Event(), threading.Event()
        self._paused = False
        if start:
            self.start()

    # ---- public ---------------------------------------------------------------------------------
    @property
    def global_epoch(self) -> int:
        return self.global_progress.epoch

    @property
    def ready_to_update_epoch(self) -> bool:
        return (self.global_epoch > self.local_progress.epoch
                or self.global_progress.samples_accumulated >= self.target_batch_size
                or get_dht_time() >= self.global_progress.eta_next_epoch)

    @property
    def estimated_next_update_time(self) -> float:
        return get_dht_time() if self.ready_to_update_epoch else self.global_progress.eta_next_epoch

    def _local(self, epoch: int, samples: int) -> LocalTrainingProgress:
        return LocalTrainingProgress(peer_id=self.dht.peer_id.raw, epoch=epoch, samples_accumulated=samples,
                                     samples_per_second=float(self.performance_ema.samples_per_second),
                                     time=get_dht_time(), client_mode=self.client_mode)

    def report_local_progress(self, local_epoch: int, samples_accumulated: int, update_global_samples: bool = True):
        extra = samples_accumulated - self.local_progress.samples_accumulated
        if update_global_samples and local_epoch == self.local_progress.epoch == self.global_progress.epoch:
            self.global_progress.samples_accumulated += extra
        if extra > 0:
            self.performance_ema.update(task_size=extra)
        else:
            self.performance_ema.reset_timer()
        with self.lock_local:
            self.local_progress = self._local(local_epoch, samples_accumulated)
        self.should_report.set()

    @contextlib.contextmanager
    def pause_updates(self):
        with self.lock_global, self.performance_ema.pause():
            self._paused = True
            try:
                yield
            finally:
                self._paused = False

    def update_epoch(self, new_epoch: Optional[int] = None) -> int:
        if new_epoch is None:
            new_epoch = self.local_progress.epoch + 1
        if new_epoch > self.global_progress.epoch:
            self.global_progress.epoch = new_epoch
            self.global_progress.samples_accumulated = 0
            self.global_progress.eta_next_epoch = float("inf")
        self.report_local_progress(new_epoch, samples_accumulated=0)
        self.fetched.clear()
        self._fetch_now.set()
        return new_epoch

    # ---- background loops -------------------------------------------------------------------------
    def run(self):
        self._fetch_now = threading.Event()
        reporter = threading.Thread(target=self._reporter, daemon=True)
        reporter.start()
        try:
            while not self.shutdown_triggered.is_set():
                wait = max(0.0, self.global_progress.next_fetch_time - get_dht_time())
                self._fetch_now.wait(wait)
                self._fetch_now.clear()
                if self.shutdown_triggered.is_set():
                    break
                try:
                    self._fetch()
                except Exception:
                    time.sleep(self.min_refresh_period)
        finally:
            self.shutdown_complete.set()

    def _reporter(self):
        last = -float("inf")
        while not self.shutdown_triggered.is_set():
            self.should_report.wait(self.metadata_expiration / 2)
            self.should_report.clear()
            if self.shutdown_triggered.is_set():
                return
            wait = self.min_refresh_period - (get_dht_time() - last)
            if wait > 0:
                time.sleep(wait)
            with self.lock_local:
                rec = self.local_progress.model_dump()
            try:
                self.dht.store(self.training_progress_key, rec, get_dht_time() + self.metadata_expiration,
                               subkey=self._local_public_key)
                last = get_dht_time()
            except Exception:
                pass

    def _fetch(self):
        entry = self.dht.get(self.training_progress_key, latest=True)
        records = []
        if entry is not None and isinstance(entry.value, dict):
            for v in entry.value.values():
                if v.value is not None:
                    try:
                        records.append(LocalTrainingProgress.model_validate(v.value))
                    except Exception:
                        pass
        with self.lock_global:
            self.global_progress = self._aggregate(records)
        self.fetched.set()

    def _aggregate(self, records) -> GlobalTrainingProgress:
        now = get_dht_time()
        with self.lock_local:
            local = self.local_progress
        records = [r for r in records if r.peer_id != local.peer_id] + [local]
        epoch = max(r.epoch for r in records)
        total, perf, peers, clients = 0, 0.0, 0, 0
        for r in records:
            peers += 1
            clients += int(r.client_mode)
            if r.epoch == epoch:
                # extrapolate by the peer's own rate for the age of its record (bounded by one refresh period)
                age = min(max(0.0, now - r.time), self.max_refresh_period)
                total += r.samples_accumulated + (int(age * r.samples_per_second) if r.peer_id != local.peer_id else 0)
                perf += r.samples_per_second
            elif r.epoch == epoch - 1:
                perf += r.samples_per_second
        remaining = max(0, self.target_batch_size - total)
        eta = now + remaining / max(perf, 1e-9) if remaining > 0 else now
        if perf <= 1e-9 and remaining > 0:
            eta = float("inf")
        drift = max(0.0, eta - now) / max(1.0, self.expected_drift_peers) if eta != float("inf") else self.default_refresh_period
        period = min(self.max_refresh_period, max(self.min_refresh_period, drift * self.expected_drift_rate + self.min_refresh_period))
        return GlobalTrainingProgress(epoch, total, self.target_batch_size, peers, clients, eta, now + period)

    def shutdown(self, timeout: Optional[float] = 2.0):
        self.shutdown_triggered.set()
        self.should_report.set()
        if hasattr(self, "_fetch_now"):
            self._fetch_now.set()
        self.shutdown_complete.wait(timeout)
        try:
            self.dht.store(self.training_progress_key, None, get_dht_time() + self.metadata_expiration,
                           subkey=self._local_public_key)
        except Exception:
            pass
