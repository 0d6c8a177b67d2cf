"""Checkpoint storage back-ends.

Reference: ``src/neuronx_distributed/trainer/checkpoint_storage.py:28-559`` -- ``BaseCheckpointStorage`` API,
``FilesysCheckpointStorage`` and ``S3CheckpointStorage`` (boto3 multipart, retry with exponential back-off,
slow-down detection), factory by ``s3://`` prefix.
"""
from __future__ import annotations

import io
import os
import random
import re
import shutil
import time
from typing import Tuple, Any, List, Optional

import torch


class BaseCheckpointStorage:
    def __init__(self, dirname: str):
        self._dirname = dirname

    def dirname(self) -> str:
        return self._dirname

    # --- abstract API (reference :28-117) ---
    def dir_exists(self, dirname: str) -> bool: raise NotImplementedError
    def file_exists(self, filename: str) -> bool: raise NotImplementedError
    def is_checkpoint_xser(self, dirname: str) -> bool: raise NotImplementedError
    def list_checkpoint_tags(self) -> List[str]: raise NotImplementedError
    def list_completed_checkpoint_tags(self) -> List[str]: raise NotImplementedError
    def find_files(self, dirname: str, pattern: str) -> List[str]: raise NotImplementedError
    def find_subdirs_contain_path(self, pattern: str, search_depth: int = 1) -> List[str]: raise NotImplementedError
    def save_text(self, text: str, filename: str) -> None: raise NotImplementedError
    def save_object(self, obj: Any, filename: str) -> None: raise NotImplementedError
    def load_object(self, filename: str, map_location=None) -> Any: raise NotImplementedError
    def create_dir(self, dirname: str, exist_ok: bool = True) -> None: raise NotImplementedError
    def create_shared_dir(self, dirname: str, exist_ok: bool = True, process_group=None) -> None: raise NotImplementedError
    def remove_dir(self, dirname: str) -> None: raise NotImplementedError
    def remove_file(self, filename: str) -> None: raise NotImplementedError
    def remove_dirs(self, dirnames: List[str]) -> None:
        for d in dirnames:
            self.remove_dir(d)
    def remove_files(self, filenames: List[str]) -> None:
        for f in filenames:
            self.remove_file(f)


class FilesysCheckpointStorage(BaseCheckpointStorage):
    """Plain file system (reference :120-216).  Writes go to ``<name>.tmp.<pid>`` then ``os.replace`` so a crash
    mid-write never leaves a truncated file under the final name."""

    def _p(self, name: str) -> str:
        return os.path.join(self._dirname, name)

    def dir_exists(self, dirname: str) -> bool:
        return os.path.isdir(self._p(dirname))

    def file_exists(self, filename: str) -> bool:
        return os.path.isfile(self._p(filename))

    def is_checkpoint_xser(self, dirname: str) -> bool:
        root = self._p(dirname)
        for sub in ("model", "optim"):
            d = os.path.join(root, sub)
            if os.path.isdir(d) and any(e.endswith(".tensors") for e in os.listdir(d)):
                return True
        return False

    def _tags_with(self, marker: str) -> List[str]:
        if not os.path.isdir(self._dirname):
            return []
        tags = [t for t in os.listdir(self._dirname) if os.path.exists(os.path.join(self._dirname, t, marker))]
        tags.sort(key=lambda t: os.path.getmtime(os.path.join(self._dirname, t, marker)))
        return tags

    def list_checkpoint_tags(self) -> List[str]:
        return self._tags_with("checkpoint")

    def list_completed_checkpoint_tags(self) -> List[str]:
        return self._tags_with("done")

    def find_files(self, dirname: str, pattern: str) -> List[str]:
        rx = re.compile(pattern)
        out = []
        root = self._p(dirname)
        for base, _, files in os.walk(root):
            for f in files:
                rel = os.path.relpath(os.path.join(base, f), self._dirname)
                if rx.search(rel):
                    out.append(rel)
        return sorted(out)

    def find_subdirs_contain_path(self, pattern: str, search_depth: int = 1) -> List[str]:
        return [t for t in (os.listdir(self._dirname) if os.path.isdir(self._dirname) else [])
                if os.path.exists(os.path.join(self._dirname, t, pattern))]

    def save_text(self, text: str, filename: str) -> None:
        path = self._p(filename)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        tmp = f"{path}.tmp.{os.getpid()}"
        with open(tmp, "w") as f:
            f.write(text)
        os.replace(tmp, path)

    def save_object(self, obj: Any, filename: str) -> None:
        path = self._p(filename)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        tmp = f"{path}.tmp.{os.getpid()}"
        torch.save(obj, tmp)
        os.replace(tmp, path)

    def load_object(self, filename: str, map_location=None) -> Any:
        return torch.load(self._p(filename), map_location=map_location, weights_only=False)

    def create_dir(self, dirname: str, exist_ok: bool = True) -> None:
        os.makedirs(self._p(dirname), exist_ok=exist_ok)

    def create_shared_dir(self, dirname: str, exist_ok: bool = True, process_group=None) -> None:
        os.makedirs(self._p(dirname), exist_ok=True)

    def remove_dir(self, dirname: str) -> None:
        shutil.rmtree(self._p(dirname), ignore_errors=True)

    def remove_file(self, filename: str) -> None:
        try:
            os.remove(self._p(filename))
        except FileNotFoundError:
            pass


class S3CheckpointStorage(BaseCheckpointStorage):
    """S3 back-end (reference :219-555): multipart upload in 64 MiB parts, retries with exponential back-off
    and jitter (up to 12 attempts), "SlowDown"/throttling detection.  ``boto3`` is imported lazily; the class is
    constructible without it (tests mock the client)."""

    PART_SIZE = 64 * 1024 * 1024
    MAX_ATTEMPTS = 12

    def __init__(self, dirname: str, client=None):
        super().__init__(dirname)
        m = re.match(r"s3://([^/]+)/?(.*)", dirname)
        if not m:
            raise ValueError(f"not an s3 url: {dirname}")
        self._bucket, self._base_key = m.group(1), m.group(2).rstrip("/")
        self._client = client

    @property
    def client(self):
        if self._client is None:
            import boto3  # noqa: WPS433 (lazy: not installed on the training image)

            self._client = boto3.session.Session().client("s3")
        return self._client

    # -- reference method names (``checkpoint_storage.py:219-555``) ---------------------------------------------------
    @staticmethod
    def parse_path(path: str) -> Tuple[str, str]:
        """``s3://bucket/key/...`` -> ``(bucket, key)``."""
        m = re.match(r"s3://([^/]+)/?(.*)", path)
        if not m:
            raise ValueError(f"not an s3 url: {path}")
        return m.group(1), m.group(2)

    def convert_path_to_key(self, path: str) -> str:
        return self._k(path)

    def get_client(self):
        return self.client

    def get_resource(self):
        import boto3  # noqa: WPS433

        return boto3.session.Session().resource("s3")

    @staticmethod
    def is_slow_down_error(exc: Exception) -> bool:
        return S3CheckpointStorage._is_slow_down(exc)

    def s3_action_with_retry(self, fn, *a, **kw):
        return self._retry(fn, *a, **kw)

    def upload_stream_to_file(self, stream, filename: str) -> None:
        self._upload(stream.getvalue() if hasattr(stream, "getvalue") else stream.read(), self._k(filename))

    def download_file_to_stream(self, filename: str, stream) -> None:
        resp = self._retry(self.client.get_object, Bucket=self._bucket, Key=self._k(filename))
        stream.write(resp["Body"].read())
        stream.seek(0)

    def _k(self, name: str) -> str:
        return f"{self._base_key}/{name}".strip("/") if name else self._base_key

    @staticmethod
    def _is_slow_down(exc: Exception) -> bool:
        text = str(exc)
        return any(s in text for s in ("SlowDown", "Throttling", "RequestLimitExceeded", "503"))

    def _retry(self, fn, *a, **kw):
        delay = 0.5
        for attempt in range(self.MAX_ATTEMPTS):
            try:
                return fn(*a, **kw)
            except Exception as e:  # noqa: BLE001
                if attempt == self.MAX_ATTEMPTS - 1:
                    raise
                factor = 4.0 if self._is_slow_down(e) else 2.0
                time.sleep(min(delay, 60.0) * (1.0 + random.random() * 0.25))
                delay *= factor

    def _list(self, prefix: str, delimiter: Optional[str] = None):
        kw = dict(Bucket=self._bucket, Prefix=prefix)
        if delimiter:
            kw["Delimiter"] = delimiter
        keys, prefixes, token = [], [], None
        while True:
            if token:
                kw["ContinuationToken"] = token
            resp = self._retry(self.client.list_objects_v2, **kw)
            keys += [(o["Key"], o.get("LastModified")) for o in resp.get("Contents", [])]
            prefixes += [p["Prefix"] for p in resp.get("CommonPrefixes", [])]
            if not resp.get("IsTruncated"):
                break
            token = resp.get("NextContinuationToken")
        return keys, prefixes

    def dir_exists(self, dirname: str) -> bool:
        keys, prefixes = self._list(self._k(dirname).rstrip("/") + "/", "/")
        return bool(keys or prefixes)

    def file_exists(self, filename: str) -> bool:
        keys, _ = self._list(self._k(filename))
        return any(k == self._k(filename) for k, _ in keys)

    def is_checkpoint_xser(self, dirname: str) -> bool:
        keys, _ = self._list(self._k(dirname).rstrip("/") + "/")
        return any(".tensors/" in k for k, _ in keys)

    def _tags_with(self, marker: str) -> List[str]:
        base = (self._base_key + "/") if self._base_key else ""
        _, prefixes = self._list(base, "/")
        found = []
        for p in prefixes:
            tag = p[len(base):].rstrip("/")
            keys, _ = self._list(f"{base}{tag}/{marker}")
            for k, ts in keys:
                if k == f"{base}{tag}/{marker}":
                    found.append((ts, tag))
        found.sort(key=lambda x: (x[0] is None, x[0]))
        return [t for _, t in found]

    def list_checkpoint_tags(self) -> List[str]:
        return self._tags_with("checkpoint")

    def list_completed_checkpoint_tags(self) -> List[str]:
        return self._tags_with("done")

    def find_files(self, dirname: str, pattern: str) -> List[str]:
        rx = re.compile(pattern)
        base = (self._base_key + "/") if self._base_key else ""
        keys, _ = self._list(self._k(dirname).rstrip("/") + "/")
        return sorted(k[len(base):] for k, _ in keys if rx.search(k[len(base):]))

    def find_subdirs_contain_path(self, pattern: str, search_depth: int = 1) -> List[str]:
        return self._tags_with(pattern)

    def _upload(self, data: bytes, key: str) -> None:
        if len(data) <= self.PART_SIZE:
            self._retry(self.client.put_object, Bucket=self._bucket, Key=key, Body=data)
            return
        mp = self._retry(self.client.create_multipart_upload, Bucket=self._bucket, Key=key)
        parts = []
        try:
            for i, off in enumerate(range(0, len(data), self.PART_SIZE), start=1):
                r = self._retry(self.client.upload_part, Bucket=self._bucket, Key=key, PartNumber=i,
                                UploadId=mp["UploadId"], Body=data[off:off + self.PART_SIZE])
                parts.append({"ETag": r["ETag"], "PartNumber": i})
            self._retry(self.client.complete_multipart_upload, Bucket=self._bucket, Key=key,
                        UploadId=mp["UploadId"], MultipartUpload={"Parts": parts})
        except Exception:
            self.client.abort_multipart_upload(Bucket=self._bucket, Key=key, UploadId=mp["UploadId"])
            raise

    def save_text(self, text: str, filename: str) -> None:
        self._upload(text.encode(), self._k(filename))

    def save_object(self, obj: Any, filename: str) -> None:
        buf = io.BytesIO()
        torch.save(obj, buf)
        self._upload(buf.getvalue(), self._k(filename))

    def load_object(self, filename: str, map_location=None) -> Any:
        resp = self._retry(self.client.get_object, Bucket=self._bucket, Key=self._k(filename))
        return torch.load(io.BytesIO(resp["Body"].read()), map_location=map_location, weights_only=False)

    def create_dir(self, dirname: str, exist_ok: bool = True) -> None:
        pass  # S3 has no directories

    def create_shared_dir(self, dirname: str, exist_ok: bool = True, process_group=None) -> None:
        pass

    def remove_dir(self, dirname: str) -> None:
        keys, _ = self._list(self._k(dirname).rstrip("/") + "/")
        for i in range(0, len(keys), 1000):
            self._retry(self.client.delete_objects, Bucket=self._bucket,
                        Delete={"Objects": [{"Key": k} for k, _ in keys[i:i + 1000]]})

    def remove_file(self, filename: str) -> None:
        self._retry(self.client.delete_object, Bucket=self._bucket, Key=self._k(filename))


def create_checkpoint_storage(dirname: str) -> BaseCheckpointStorage:
    """``s3://...`` -> S3, anything else -> file system (reference :558-559)."""
    return S3CheckpointStorage(dirname) if dirname.startswith("s3://") else FilesysCheckpointStorage(dirname)
