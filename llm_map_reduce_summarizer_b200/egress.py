"""Egress formats of the path.

`save_chunk_records` writes the `--save-chunks` file of the reference's CLI (/root/reference/main.py:178-201, format
README.md:145-158) from the chunk dicts this package returns.  What comes from the device buffers is `chunk_index`,
`start_time` and `end_time` (TRC_A_CHUNKS records; the times are the caller's own Python numbers, an int stays an int, so
their JSON spelling is Python's and is produced here, on the host); `summary` and `tokens_used` are filled in by the map
step (llm_executor.py, host side, out of scope) and default to "" and 0 exactly as main.py's `chunk.get(...)` does.
"""
import datetime
import json
from typing import Any, Dict, List, Optional


def chunk_records(processed_chunks: List[Dict[Any, Any]], timestamp: Optional[str] = None) -> Dict[str, Any]:
    """The object main.py:180-194 builds."""
    return {"timestamp": timestamp if timestamp is not None else datetime.datetime.now().strftime("%Y-%m-%d %H:%M:%S"),
            "chunks": [{"chunk_index": c.get("chunk_index", -1), "start_time": c.get("start_time", ""),
                        "end_time": c.get("end_time", ""), "summary": c.get("summary", ""),
                        "tokens_used": c.get("tokens_used", 0)} for c in processed_chunks]}


def save_chunk_records(processed_chunks: List[Dict[Any, Any]], path: str, timestamp: Optional[str] = None) -> None:
    """main.py:196-197: json.dump(..., indent=2) in UTF-8."""
    with open(path, "w", encoding="utf-8") as f:
        json.dump(chunk_records(processed_chunks, timestamp), f, indent=2)
