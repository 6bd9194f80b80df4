"""Id and path types of the object-store boundary.

Same two patterns as the reference's ``utils/validation.py:19-22`` (and the
protovalidate rules at ``code_interpreter_service.proto:30-33``): a SHA-256 hex
digest (64 lowercase hex characters) satisfies ``Hash``.
"""
from typing import Annotated, TypeAliasType

from pydantic import Field

HASH_PATTERN = r"^[0-9a-zA-Z_-]{1,255}$"
ABSOLUTE_PATH_PATTERN = r"^/[^/].*$"

Hash = TypeAliasType("Hash", Annotated[str, Field(pattern=HASH_PATTERN)])
AbsolutePath = TypeAliasType("AbsolutePath", Annotated[str, Field(pattern=ABSOLUTE_PATH_PATTERN)])
