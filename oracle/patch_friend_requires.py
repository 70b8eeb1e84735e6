#!/usr/bin/env python3
"""TEST INFRASTRUCTURE ONLY (oracle/_ref build).

Regenerates one seqan3 header with the `requires ...` clause removed from *friend class template declarations*
(`template <...> requires ... friend class X;`).  GCC 13 rejects those as "redeclaration with different
constraints"; dropping the constraint on a friend declaration does not change any computed result
(SURVEY.md §8(c) step 1).  Input is read from the reference tree, output is written under oracle/_ref/shim/
(git-ignored) — nothing is copied into the repository.

usage: patch_friend_requires.py <in.hpp> <out.hpp>
"""
import re
import sys


def patch(text: str) -> tuple[str, int]:
    # template <...>\n   requires <anything not containing ';'>\n   friend class NAME;
    pat = re.compile(r"(template\s*<[^;{}]*?>)\s*requires\b[^;{}]*?(\bfriend\s+class\s+\w+\s*;)", re.S)
    return pat.subn(lambda m: m.group(1) + "\n    " + m.group(2), text)


def main() -> int:
    src, dst = sys.argv[1], sys.argv[2]
    with open(src, "r", encoding="utf-8") as fh:
        text = fh.read()
    out, n = patch(text)
    if n == 0:
        sys.stderr.write(f"patch_friend_requires: no constrained friend declaration found in {src}\n")
        return 1
    with open(dst, "w", encoding="utf-8") as fh:
        fh.write(out)
    return 0


if __name__ == "__main__":
    sys.exit(main())
