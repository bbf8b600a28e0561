"""Generate tests/golden/cache_prompts.json (AUTHORING container only; needs /root/reference).

The keys of the reference's .cache/42/cache.db are the pickled LLM requests, i.e. the prompts
RFQFieldGenerator._generate_sync built.  They are extracted WITHOUT unpickling (pickletools.genops,
longest str operand) and kept when they start with the reference's CURRENT template, which is read
from app/rfq_agent.py with `ast` (the module cannot be imported: autogen is absent).
    python tests/golden/make_prompt_goldens.py
"""
import ast
import json
import pickletools
import sqlite3
from pathlib import Path

REF = Path("/root/reference")
HERE = Path(__file__).resolve().parent


def main():
    template = None
    for node in ast.parse((REF / "app/rfq_agent.py").read_text()).body:
        if isinstance(node, ast.Assign) and getattr(node.targets[0], "id", None) == "EXTRACTION_PROMPT_TEMPLATE":
            template = ast.literal_eval(node.value)
    con = sqlite3.connect(f"file:{REF}/.cache/42/cache.db?mode=ro&immutable=1", uri=True)
    out = {"template_sha256": __import__("hashlib").sha256(template.encode()).hexdigest(), "prompts": {}}
    for rid, key in con.execute("select rowid, key from Cache order by rowid"):
        prompt = max((a for _, a, _ in pickletools.genops(bytes(key)) if isinstance(a, str)), key=len)
        if prompt.startswith(template):
            out["prompts"][str(rid)] = prompt
    (HERE / "cache_prompts.json").write_text(json.dumps(out, ensure_ascii=False, indent=0))
    # the values of the same table are the LLM's replies to those requests: what
    # RFQFieldGenerator._parse_and_validate_response received (app/rfq_agent.py:185-190)
    replies = {}
    for rid, val in con.execute("select rowid, value from Cache order by rowid"):
        replies[str(rid)] = max((a for _, a, _ in pickletools.genops(bytes(val)) if isinstance(a, str)), key=len)
    (HERE / "cache_replies.json").write_text(json.dumps(replies, ensure_ascii=False, indent=0))
    print({k: len(v) for k, v in out["prompts"].items()})


if __name__ == "__main__":
    main()
