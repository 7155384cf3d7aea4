"""Hugging Face based authorization (reference huggingface_auth.py:19-193).

The authority is an HTTPS endpoint: ``PUT {auth}/api/experiments/join`` with the user's HF access token and this
peer's RSA public key returns the authority's public key, the coordinator address and a signed ``AccessToken``.
There is no network on the build/benchmark hosts: ``--authorize false`` skips all of this (and, unlike the reference,
does not crash: task.py falls back to a host-derived run name).
"""
from __future__ import annotations

import base64
import os
import time
from datetime import datetime, timedelta
from getpass import getpass
from typing import Optional

from dalle_b200.collab.auth import AccessToken, TokenAuthorizerBase
from dalle_b200.dht.crypto import RSAPrivateKey, RSAPublicKey


class NonRetriableError(Exception):
    pass


def call_with_retries(func, n_retries: int = 10, initial_delay: float = 1.0):
    """Exponential back-off; ``NonRetriableError`` aborts at once."""
    for attempt in range(n_retries):
        try:
            return func()
        except NonRetriableError:
            raise
        except Exception:
            if attempt == n_retries - 1:
                raise
            time.sleep(initial_delay * (2 ** attempt))


class InvalidCredentialsError(NonRetriableError):
    pass


class NotInAllowlistError(NonRetriableError):
    pass


class HuggingFaceAuthorizer(TokenAuthorizerBase):
    _AUTH_SERVER_URL = "https://collaborative-training-auth.huggingface.co"

    def __init__(self, organization_name: str, model_name: str, hf_user_access_token: str):
        super().__init__()
        self.organization_name, self.model_name = organization_name, model_name
        self.hf_user_access_token = hf_user_access_token
        self._authority_public_key: Optional[RSAPublicKey] = None
        self.coordinator_ip = self.coordinator_port = None
        try:
            import huggingface_hub

            self.username = huggingface_hub.HfApi().whoami(token=hf_user_access_token)["name"]
        except Exception:
            self.username = os.environ.get("HF_USERNAME", "unknown")

    async def get_token(self) -> AccessToken:
        return call_with_retries(self._join_experiment)

    def _join_experiment(self) -> AccessToken:
        import requests

        url = f"{self._AUTH_SERVER_URL}/api/experiments/join"
        headers = {"Authorization": f"Bearer {self.hf_user_access_token}"}
        response = requests.put(url, headers=headers, timeout=30, params={
            "organization_name": self.organization_name, "model_name": self.model_name},
            json={"experiment_join_input": {"peer_public_key": self.local_public_key.to_bytes().decode()}})
        if response.status_code == 401:
            raise InvalidCredentialsError()
        if response.status_code == 403:
            raise NotInAllowlistError()
        response.raise_for_status()
        body = response.json()
        self._authority_public_key = RSAPublicKey.from_bytes(body["auth_server_public_key"].encode())
        self.coordinator_ip, self.coordinator_port = body["coordinator_ip"], body["coordinator_port"]
        tok = body["hivemind_access"]
        return AccessToken(username=tok["username"], public_key=tok["peer_public_key"].encode(),
                           expiration_time=str(datetime.fromisoformat(tok["expiration_time"])),
                           signature=tok["signature"].encode())

    @staticmethod
    def _token_to_bytes(token: AccessToken) -> bytes:
        return f"{token.username} {token.public_key} {token.expiration_time}".encode()

    def is_token_valid(self, access_token: AccessToken) -> bool:
        if self._authority_public_key is None:
            return False
        if not self._authority_public_key.verify(self._token_to_bytes(access_token), access_token.signature):
            return False
        try:
            expires = datetime.fromisoformat(access_token.expiration_time)
        except ValueError:
            return False
        return expires >= datetime.utcnow()  # naive UTC, as the authority issues it

    _MAX_LATENCY = timedelta(minutes=1)

    def does_token_need_refreshing(self, access_token: AccessToken) -> bool:
        return datetime.fromisoformat(access_token.expiration_time) < datetime.utcnow() + self._MAX_LATENCY


def authorize_with_huggingface() -> HuggingFaceAuthorizer:
    """Credentials from HF_ORGANIZATION_NAME / HF_MODEL_NAME / HF_USER_ACCESS_TOKEN, prompting when interactive."""
    while True:
        org = os.getenv("HF_ORGANIZATION_NAME") or input("HuggingFace organization name: ")
        model = os.getenv("HF_MODEL_NAME") or input("HuggingFace model name: ")
        token = os.getenv("HF_USER_ACCESS_TOKEN")
        if token is None:
            try:
                from huggingface_hub import HfFolder

                token = HfFolder.get_token()
            except Exception:
                token = None
        if token is None:
            token = getpass("HuggingFace user access token (characters will be hidden): ")
        authorizer = HuggingFaceAuthorizer(org, model, token)
        try:
            import asyncio

            asyncio.run(authorizer.refresh_token_if_needed())
            return authorizer
        except InvalidCredentialsError:
            print("Invalid user access token, please try again")
            if os.getenv("HF_USER_ACCESS_TOKEN"):
                raise
        except NotInAllowlistError:
            print("This account is not specified by organization admins in the allowlist")
            raise
