"""RSA keys for record signatures and access tokens (replaces hivemind.utils.crypto)."""
from __future__ import annotations

import base64
import threading
from typing import Optional

from cryptography import exceptions
from cryptography.hazmat.primitives import hashes, serialization
from cryptography.hazmat.primitives.asymmetric import padding, rsa


class RSAPublicKey:
    def __init__(self, key):
        self._key = key

    def verify(self, data: bytes, signature: bytes) -> bool:
        try:
            self._key.verify(base64.b64decode(signature), data, _PADDING, hashes.SHA256())
            return True
        except (ValueError, exceptions.InvalidSignature):
            return False

    def to_bytes(self) -> bytes:
        return self._key.public_bytes(encoding=serialization.Encoding.OpenSSH,
                                      format=serialization.PublicFormat.OpenSSH)

    @classmethod
    def from_bytes(cls, key: bytes) -> "RSAPublicKey":
        k = serialization.load_ssh_public_key(key)
        if not isinstance(k, rsa.RSAPublicKey):
            raise ValueError(f"Expected an RSA public key, got {k}")
        return cls(k)


_PADDING = padding.PSS(mgf=padding.MGF1(hashes.SHA256()), salt_length=padding.PSS.MAX_LENGTH)


class RSAPrivateKey:
    _process_wide_key: Optional["RSAPrivateKey"] = None
    _lock = threading.Lock()

    def __init__(self, key=None):
        self._key = key or rsa.generate_private_key(public_exponent=65537, key_size=2048)

    @classmethod
    def process_wide(cls) -> "RSAPrivateKey":
        with cls._lock:
            if cls._process_wide_key is None:
                cls._process_wide_key = cls()
            return cls._process_wide_key

    def sign(self, data: bytes) -> bytes:
        return base64.b64encode(self._key.sign(data, _PADDING, hashes.SHA256()))

    def get_public_key(self) -> RSAPublicKey:
        return RSAPublicKey(self._key.public_key())

    def to_bytes(self) -> bytes:
        return self._key.private_bytes(serialization.Encoding.PEM, serialization.PrivateFormat.PKCS8,
                                       serialization.NoEncryption())

    @classmethod
    def from_bytes(cls, data: bytes) -> "RSAPrivateKey":
        return cls(serialization.load_pem_private_key(data, password=None))
