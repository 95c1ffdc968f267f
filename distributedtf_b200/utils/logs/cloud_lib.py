"""Cloud environment probe.  The reference does an HTTP GET to the GCE metadata host on every run
(``logs/cloud_lib.py:21-34`` via ``logger.py:419-422``); on a machine without network that is a 5 s
timeout per member per round.  Here the probe is opt-in (``DTF_ALLOW_NETWORK_PROBE=1``)."""
import os

GCP_METADATA_URL = "http://metadata/computeMetadata/v1/instance/hostname"
GCP_METADATA_HEADER = {"Metadata-Flavor": "Google"}


def on_gcp() -> bool:
    if os.environ.get("DTF_ALLOW_NETWORK_PROBE") != "1":
        return False
    try:
        import requests
        return requests.get(GCP_METADATA_URL, headers=GCP_METADATA_HEADER, timeout=5).status_code == 200
    except Exception:
        return False
