"""DHT record validators: RSA-signed sub-keys and pydantic schemas.

* ``RSASignatureValidator``: a key or sub-key containing ``[owner:<ssh-rsa public key>]`` may only be written with a
  value carrying a valid ``[signature:<b64>]`` suffix made by that key over (key, subkey, value, expiration);
  the suffix is stripped on read.  Used for ``{prefix}_metrics`` (reference utils.py:27-30, callback.py:80-86).
* ``SchemaValidator``: for a pydantic model with fields f1, f2..., DHT key ``{prefix}_{field}`` must hold a value
  of the field's type; ``Dict[BytesWithPublicKey, X]`` fields validate sub-keys/values individually.
"""
from __future__ import annotations

import dataclasses
import hashlib
import re
from typing import Any, Dict, Iterable, List, Optional

import msgpack
import pydantic

from .crypto import RSAPrivateKey, RSAPublicKey


@dataclasses.dataclass(frozen=True)
class DHTRecord:
    key: bytes          # raw (un-hashed) key bytes
    subkey: bytes       # serialized sub-key, or IS_REGULAR_VALUE
    value: bytes        # serialized value
    expiration_time: float


IS_REGULAR_VALUE = b""


def pack(obj) -> bytes:
    return msgpack.packb(obj, use_bin_type=True)


def unpack(data: bytes):
    return msgpack.unpackb(data, raw=False, strict_map_key=False)


class RecordValidatorBase:
    def validate(self, record: DHTRecord) -> bool:
        return True

    def sign_value(self, record: DHTRecord) -> bytes:
        return record.value

    def strip_value(self, record: DHTRecord) -> bytes:
        return record.value

    @property
    def priority(self) -> int:
        return 0

    def merge_with(self, other: "RecordValidatorBase") -> bool:
        return False


class CompositeValidator(RecordValidatorBase):
    def __init__(self, validators: Iterable[RecordValidatorBase] = ()):
        self._validators: List[RecordValidatorBase] = []
        self.extend(validators)

    def extend(self, validators: Iterable[RecordValidatorBase]):
        for v in validators:
            if not any(type(v) is type(o) and o.merge_with(v) for o in self._validators):
                self._validators.append(v)
        self._validators.sort(key=lambda v: -v.priority)

    def validate(self, record: DHTRecord) -> bool:
        # outermost (highest priority, e.g. the signature) first; each level strips its envelope for the next
        for i, v in enumerate(self._validators):
            if not v.validate(record):
                return False
            if i < len(self._validators) - 1:
                record = dataclasses.replace(record, value=v.strip_value(record))
        return True

    def sign_value(self, record: DHTRecord) -> bytes:
        for v in reversed(self._validators):  # innermost first, signature last
            record = dataclasses.replace(record, value=v.sign_value(record))
        return record.value

    def strip_value(self, record: DHTRecord) -> bytes:
        for v in self._validators:
            record = dataclasses.replace(record, value=v.strip_value(record))
        return record.value


class RSASignatureValidator(RecordValidatorBase):
    PUBLIC_KEY_FORMAT = b"[owner:_key_]"
    SIGNATURE_FORMAT = b"[signature:_value_]"
    PUBLIC_KEY_REGEX = re.escape(PUBLIC_KEY_FORMAT).replace(b"_key_", rb"(.+?)")
    _PUBLIC_KEY_RE = re.compile(PUBLIC_KEY_REGEX)
    _SIGNATURE_RE = re.compile(re.escape(SIGNATURE_FORMAT).replace(b"_value_", rb"(.+?)"))

    def __init__(self, private_key: Optional[RSAPrivateKey] = None):
        self._private_key = private_key or RSAPrivateKey.process_wide()
        serialized = self._private_key.get_public_key().to_bytes()
        self._local_public_key = self.PUBLIC_KEY_FORMAT.replace(b"_key_", serialized)

    @property
    def local_public_key(self) -> bytes:
        return self._local_public_key

    def _owners(self, record: DHTRecord):
        return self._PUBLIC_KEY_RE.findall(record.key) + self._PUBLIC_KEY_RE.findall(record.subkey)

    def validate(self, record: DHTRecord) -> bool:
        owners = self._owners(record)
        if not owners:
            return True
        if len(set(owners)) > 1:
            return False
        sigs = self._SIGNATURE_RE.findall(record.value)
        if len(sigs) != 1:
            return False
        try:
            key = RSAPublicKey.from_bytes(owners[0])
        except Exception:
            return False
        stripped = dataclasses.replace(record, value=self.strip_value(record))
        return key.verify(self._serialize(stripped), sigs[0])

    def sign_value(self, record: DHTRecord) -> bytes:
        if self._local_public_key not in record.key and self._local_public_key not in record.subkey:
            return record.value
        sig = self._private_key.sign(self._serialize(record))
        return record.value + self.SIGNATURE_FORMAT.replace(b"_value_", sig)

    def strip_value(self, record: DHTRecord) -> bytes:
        return self._SIGNATURE_RE.sub(b"", record.value)

    @staticmethod
    def _serialize(record: DHTRecord) -> bytes:
        return pack([record.key, record.subkey, record.value, float(record.expiration_time)])

    @property
    def priority(self) -> int:
        return 10  # outermost: signs the already-validated payload

    def merge_with(self, other) -> bool:
        return isinstance(other, RSASignatureValidator)


class BytesWithPublicKey(bytes):
    """Sub-key type: bytes that contain an ``[owner:...]`` marker (so the record must be signed by that owner)."""

    @classmethod
    def __get_pydantic_core_schema__(cls, source, handler):
        from pydantic_core import core_schema

        def check(v):
            if not isinstance(v, bytes):
                raise ValueError("bytes required")
            if not RSASignatureValidator._PUBLIC_KEY_RE.search(v):
                raise ValueError("subkey must contain an [owner:<public key>] marker")
            return v

        return core_schema.no_info_plain_validator_function(check)


class SchemaValidator(RecordValidatorBase):
    def __init__(self, schema: type, *, allow_extra_keys: bool = True, prefix: Optional[str] = None):
        self._schemas = [schema]
        self._prefix = prefix
        self._allow_extra = allow_extra_keys
        self._key_to_field: Dict[bytes, List[Any]] = {}
        self._register(schema, prefix)

    def _register(self, schema, prefix):
        for name, field in schema.model_fields.items():
            raw = f"{prefix}_{name}" if prefix is not None else name
            self._key_to_field.setdefault(raw.encode(), []).append((schema, name, field))

    def validate(self, record: DHTRecord) -> bool:
        fields = self._key_to_field.get(record.key)
        if fields is None:
            return self._allow_extra
        try:
            value = unpack(record.value)
        except Exception:
            return False
        for schema, name, field in fields:
            try:
                if record.subkey == IS_REGULAR_VALUE:
                    pydantic.TypeAdapter(field.annotation).validate_python(value)
                else:
                    subkey = unpack(record.subkey)
                    pydantic.TypeAdapter(field.annotation).validate_python({subkey: value})
                return True
            except pydantic.ValidationError:
                continue
        return False

    @property
    def priority(self) -> int:
        return 0

    def merge_with(self, other) -> bool:
        if not isinstance(other, SchemaValidator):
            return False
        for k, v in other._key_to_field.items():
            self._key_to_field.setdefault(k, []).extend(v)
        self._allow_extra = self._allow_extra and other._allow_extra
        return True
