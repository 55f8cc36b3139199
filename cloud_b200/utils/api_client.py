"""Local counterpart of ``TFCloudHttpRequest`` (reference: utils/google_api_client.py:21-39, which
stamps every REST call with the user agent ``tf-cloud/<version>`` for usage tracking).  There are
no REST calls here; the same tag is stamped on every record written to the local job registry and
the local Vizier store so tooling can tell which client version produced them."""
from ..version import __version__


def user_agent() -> str:
    return "cloud-b200/" + __version__


class LocalRequest:
    """A request envelope for the in-process services (job registry, local Vizier)."""

    def __init__(self, method: str, body=None, headers=None):
        self.method = method
        self.body = body
        self.headers = dict(headers or {})
        self.headers["user-agent"] = user_agent()

    def to_record(self) -> dict:
        return {"method": self.method, "body": self.body, "headers": self.headers}
