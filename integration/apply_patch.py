"""Applies integration/embedding_generator.rs.patch (line-range replacements) to a checkout of the reference and drops the new
crate in. usage: python integration/apply_patch.py /path/to/codename-symbiont [--dry-run]"""
import hashlib, os, re, shutil, sys

HERE = os.path.dirname(os.path.abspath(__file__))
WANT = "82f18c401924a791843d36906ca76a86a8c0ff21a033dd7ce0ccea6e54d4fedc"


def patched_text(original: str, patch: str) -> str:
    lines = original.split("\n")
    hunks, cur = [], None
    for l in patch.split("\n"):
        m = re.match(r"@@ (\d+)-(\d+)\s*$", l)
        if m:
            cur = [int(m.group(1)), int(m.group(2)), []]
            hunks.append(cur)
        elif cur is not None:
            cur[2].append(l)
        # lines before the first "@@" are the header
    for a, b, new in sorted(hunks, reverse=True):  # bottom up, so that earlier line numbers stay valid
        while new and new[-1] == "":
            new.pop()
        lines[a - 1:b] = new
    return "\n".join(lines)


def main():
    root, dry = sys.argv[1], "--dry-run" in sys.argv
    target = os.path.join(root, "services", "preprocessing_service", "src", "embedding_generator.rs")
    original = open(target).read()
    got = hashlib.sha256(original.encode()).hexdigest()
    if got != WANT:
        raise SystemExit(f"{target} is not the v0.3.0 file this patch was written against (sha256 {got})")
    out = patched_text(original, open(os.path.join(HERE, "embedding_generator.rs.patch")).read())
    if dry:
        sys.stdout.write(out)
        return
    open(target, "w").write(out)
    shutil.copytree(os.path.join(HERE, "libs", "b200_embed"), os.path.join(root, "libs", "b200_embed"), dirs_exist_ok=True)
    print("patched", target, "and copied libs/b200_embed; now add \"libs/b200_embed\" to the workspace members, "
          "`b200_embed = { path = \"../../libs/b200_embed\" }` and `serde = { version = \"1\", features = [\"derive\"] }` to "
          "services/preprocessing_service/Cargo.toml, and drop the three candle dependencies there")


if __name__ == "__main__":
    main()
