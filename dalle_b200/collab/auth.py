"""Request authorization (hivemind.utils.auth ``TokenAuthorizerBase`` and its ``AccessToken`` protobuf).

Every authorized RPC request carries the caller's access token (username, peer public key, expiration, all signed
by the authority), a nonce/timestamp and a signature by the caller's private key over them; servers verify the
token against the authority's public key and the request signature against the token's public key.
``HuggingFaceAuthorizer`` (huggingface_auth.py) obtains the token over HTTPS; offline runs use ``--authorize false``.
"""
from __future__ import annotations

import dataclasses
import os
import threading
import time
from datetime import datetime, timedelta
from typing import Optional

import msgpack

from ..dht.crypto import RSAPrivateKey, RSAPublicKey
from ..dht.storage import get_dht_time


@dataclasses.dataclass
class AccessToken:
    username: str = ""
    public_key: bytes = b""
    expiration_time: str = ""   # ISO format, naive UTC (reference huggingface_auth.py:116-135)
    signature: bytes = b""

    def to_dict(self):
        return dataclasses.asdict(self)

    @classmethod
    def from_dict(cls, d):
        return cls(**d)


class AuthorizerBase:
    async def sign_request_fields(self) -> dict:
        return {}

    async def validate_request_fields(self, auth: Optional[dict]) -> bool:
        return True


class TokenAuthorizerBase(AuthorizerBase):
    """Subclasses implement ``get_token``, ``is_token_valid``, ``does_token_need_refreshing``."""

    MAX_CLIENT_SERVICER_TIME_DIFF = timedelta(minutes=1)

    def __init__(self, local_private_key: Optional[RSAPrivateKey] = None):
        self._local_private_key = local_private_key or RSAPrivateKey.process_wide()
        self._local_public_key = self._local_private_key.get_public_key()
        self._local_access_token: Optional[AccessToken] = None
        self._refresh_lock = threading.Lock()
        self._recent_nonces = {}

    async def get_token(self) -> AccessToken:
        raise NotImplementedError

    def is_token_valid(self, access_token: AccessToken) -> bool:
        raise NotImplementedError

    def does_token_need_refreshing(self, access_token: AccessToken) -> bool:
        raise NotImplementedError

    async def refresh_token_if_needed(self):
        if self._local_access_token is None or self.does_token_need_refreshing(self._local_access_token):
            self._local_access_token = await self.get_token()
            assert self.is_token_valid(self._local_access_token)

    @property
    def local_public_key(self) -> RSAPublicKey:
        return self._local_public_key

    @staticmethod
    def _request_bytes(token: AccessToken, nonce: bytes, t: float) -> bytes:
        return msgpack.packb([token.to_dict(), nonce, float(t)], use_bin_type=True)

    async def sign_request_fields(self) -> dict:
        await self.refresh_token_if_needed()
        nonce, t = os.urandom(8), get_dht_time()
        sig = self._local_private_key.sign(self._request_bytes(self._local_access_token, nonce, t))
        return {"token": self._local_access_token.to_dict(), "nonce": nonce, "time": t, "sig": sig}

    async def validate_request_fields(self, auth: Optional[dict]) -> bool:
        if not auth:
            return False
        try:
            token = AccessToken.from_dict(auth["token"])
            if not self.is_token_valid(token):
                return False
            if abs(get_dht_time() - auth["time"]) > self.MAX_CLIENT_SERVICER_TIME_DIFF.total_seconds():
                return False
            key = RSAPublicKey.from_bytes(token.public_key)
            if not key.verify(self._request_bytes(token, auth["nonce"], auth["time"]), auth["sig"]):
                return False
            now = time.monotonic()
            self._recent_nonces = {n: t for n, t in self._recent_nonces.items() if now - t < 120}
            if auth["nonce"] in self._recent_nonces:
                return False  # replay
            self._recent_nonces[auth["nonce"]] = now
            return True
        except Exception:
            return False
