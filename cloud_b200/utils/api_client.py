"""Local counterpart of ``TFCloudHttpRequest`` (reference: utils/google_api_client.py:21-39, which
stamps every REST call with the user agent ``tf-cloud/<version>`` for usage tracking).  There are
no REST calls here; the same tag is stamped on every record written to the local job registry and
the local Vizier store so tooling can tell which client version produced them."""
from ..version import __version__


def user_agent() -> str:
    return "cloud-b200/" + __version__
