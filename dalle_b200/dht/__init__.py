from .dht import DHT, ValueWithExpiration, get_dht_time
from .validation import (BytesWithPublicKey, CompositeValidator, DHTRecord, RecordValidatorBase,
                         RSASignatureValidator, SchemaValidator)

__all__ = ["DHT", "ValueWithExpiration", "get_dht_time", "RecordValidatorBase", "DHTRecord", "CompositeValidator",
           "RSASignatureValidator", "SchemaValidator", "BytesWithPublicKey"]
