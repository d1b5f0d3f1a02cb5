"""Known-answer vectors for isDockerFile (internal/source/dockerfile2kube.go:117-144).

The reference has NO test for this function and the parser it calls (moby/buildkit v0.7.2,
go.mod:22) is not in the reference tree, so these answers are derived by hand from the
published parser algorithm: parity for this half is "unpinned" (DESIGN.md).  They keep the
oracle and the CUDA path honest against each other and against the written-down semantics.
Each vector: (name, bytes, is_dockerfile).
"""
L64 = b"x" * 65536
L64m = b"x" * 65535

KATS = [
    ("plain", b"FROM ubuntu\nRUN true\n", True),
    ("lowercase", b"from ubuntu:20.04\n", True),
    ("as_stage", b"FROM golang:1.15 AS build\nRUN go build\nFROM scratch\n", True),
    # the reference regexp has no \s+ after the optional --platform group, so a platform flag
    # followed by an image name does NOT match (dockerfile2kube.go:131 as written)
    ("platform", b"FROM --platform=linux/amd64 ubuntu AS x\n", False),
    ("platform_no_as", b"FROM --platform=linux/amd64 ubuntu\n", False),
    ("platform_alone", b"FROM --platform=linux/amd64\n", True),
    ("arg_first", b"ARG V=1\nARG W\nFROM ubuntu:$V\n", True),
    ("comment_first", b"# syntax=docker/dockerfile:1\n# hello\n\nFROM a\n", True),
    ("bom", b"\xef\xbb\xbfFROM a\n", True),
    ("crlf", b"FROM a\r\nRUN b\r\n", True),
    ("no_trailing_newline", b"FROM a", True),
    ("trailing_comment", b"FROM a # base image\n", True),
    ("trailing_hash_only", b"FROM a #\n", False),
    ("hash_cut", b"FROM a#b c d\n", True),
    ("two_args", b"FROM a b\n", False),
    ("second_from_matches", b"FROM a b from c\n", True),
    ("from_alone", b"FROM\n", False),
    ("from_vtab", b"FROM\x0bubuntu\n", False),
    ("leading_ws", b"   \tFROM a\n", True),
    ("leading_nbsp", b"\xc2\xa0FROM a\n", True),
    ("run_first", b"RUN echo\nFROM a\n", False),
    ("empty", b"", False),
    ("only_comments", b"# a\n# b\n", False),
    ("only_args", b"ARG A\nARG B\n", False),
    ("binary", bytes(range(256)) * 4, False),
    ("go_source", b"package main\n\nimport \"fmt\"\n", False),
    ("continuation", b"FROM \\\n  ubuntu \\\n  AS base\nRUN x\n", True),
    ("continuation_comment_inside", b"FROM \\\n# comment\n  ubuntu\n", True),
    ("continuation_empty_inside", b"FROM \\\n\n  ubuntu\n", True),
    ("continuation_eof", b"FROM ubuntu \\", True),
    # isComment() left-trims the raw continuation line first (tokenComment.Match(trimWhitespace(line))): an INDENTED
    # comment inside a continuation is dropped, it does not become part of the instruction
    ("continuation_indented_comment_dropped", b"FROM a \\\n  # c\n  b\n", False),       # "FROM a   b": two arguments
    ("continuation_indented_comment_run", b"FROM a\nRUN x \\\n    # indented\n    && y\n", True),
    ("continuation_indented_comment_env", b"FROM a\nENV A \\\n  # c\n", False),        # "ENV A": one argument
    ("continuation_nbsp_indented_comment", b"FROM a \\\n\xc2\xa0# c\n b\n", False),
    ("continuation_indented_comment_keeps_open", b"FROM \\\n \t# c \\\n  a\n", True),   # the comment's own backslash is not looked at
    ("escape_backtick", b"# escape=`\nFROM `\n  ubuntu\n", True),
    ("escape_backtick_backslash_literal", b"# escape=`\nFROM ubuntu\\\n", True),
    ("escape_invalid", b"# escape=x\nFROM a\n", False),
    ("escape_twice", b"# escape=`\n# escape=\\\nFROM a\n", False),
    ("escape_after_comment_is_comment", b"# hi\n# escape=x\nFROM a\n", True),
    ("escape_trailing_blank", b"# escape= \nFROM a\n", False),
    ("escape_no_char", b"# escape=\nFROM a\n", True),
    ("escape_case", b"#ESCAPE = `\nFROM a\n", True),
    ("env_ok_new", b"FROM a\nENV A=1 B=2\n", True),
    ("env_ok_old", b"FROM a\nENV A 1 2 3\n", True),
    ("env_one_arg", b"FROM a\nENV A\n", False),
    ("env_missing_eq", b"FROM a\nENV A=1 B\n", False),
    ("env_quoted", b"FROM a\nENV A=\"x y\" B='c d'\n", True),
    ("env_escaped_space", b"FROM a\nENV A=x\\ y\n", True),
    ("env_empty", b"FROM a\nENV\n", True),
    ("env_nbsp_only_sep", b"FROM a\nENV A\xc2\xa0B\n", False),
    ("label_missing_eq", b"FROM a\nLABEL a=b c\n", False),
    ("label_ok", b"FROM a\nLABEL \"a b\"=c\n", True),
    ("onbuild_env_bad", b"FROM a\nONBUILD ENV A\n", False),
    ("onbuild_env_ok", b"FROM a\nONBUILD ENV A=1\n", True),
    ("onbuild_nested", b"FROM a\nONBUILD ONBUILD LABEL x\n", False),
    ("onbuild_empty", b"FROM a\nONBUILD\n", True),
    ("env_with_flag", b"FROM a\nENV --foo A=1\n", True),
    ("env_flag_only", b"FROM a\nENV --foo\n", True),
    ("unknown_instruction", b"FROM a\nFOO bar\n", True),
    ("env_before_from_bad", b"ENV A\nFROM a\n", False),
    ("healthcheck", b"FROM a\nHEALTHCHECK --interval=5s CMD [\"x\"]\n", True),
    ("json_bad", b"FROM a\nCMD [\"x\n", True),
    ("long_line_65536", b"FROM a\n# " + L64[:65534] + b"\n", False),
    ("long_line_65535", b"FROM a\n# " + L64m[:65533] + b"\n", True),
    ("long_line_last_no_nl", b"FROM a\n" + L64, False),
    ("long_line_last_no_nl_ok", b"FROM a\n" + L64m, True),
    ("long_line_crlf_counts_cr", b"FROM a\n" + L64m + b"\r\n", False),
    ("from_as_long_s", b"FROM a A\xc5\xbf b\n", True),
    ("from_as_missing_name", b"FROM a AS\n", False),
    ("from_as_extra", b"FROM a AS b c\n", False),
    ("from_tab_sep", b"FROM\tubuntu\n", True),
    ("from_ff_sep", b"FROM\x0cubuntu\n", True),
    ("fromx", b"FROMX ubuntu\n", False),
    ("from_colon", b"FROM: ubuntu\n", False),
    ("nul_bytes", b"FROM a\n\x00\x00\x00\n", True),
    ("kelvin_healthcheck", b"FROM a\nHEALTHCHEC\xe2\x84\xaa NONE\n", True),
    ("dotted_I_onbuild", b"FROM a\nONBU\xc4\xb0LD ENV A\n", False),
    ("env_quote_unterminated", b"FROM a\nENV A=\"x y\n", True),
    ("env_single_quote_backslash", b"FROM a\nENV A='x\\' B\n", False),
    ("env_0x85_byte_flag_space", b"FROM a\nENV --x\x85A=1\n", True),
    ("arg_then_run", b"ARG A\nRUN x\nFROM a\n", False),
    ("whitespace_only_lines", b"   \n\t\nFROM a\n", True),
    ("backslash_only_line", b"\\\nFROM a\n", True),
    ("continuation_inside_command", b"FR\\\nOM a\n", True),
    ("continuation_inside_command_backtick", b"# escape=`\nFR`  \nOM a\nRUN x\n", True),
    ("continuation_inside_env", b"FROM a\nEN\\\nV A\n", False),
    ("backslash_not_last_in_word", b"FR\\OM a\n", False),
]


def reference_corpus():
    """(path, bytes, is_dockerfile) for the files of the reference tree recorded by tests/golden/make_dockerfile_corpus.py:
    its 16 real Dockerfiles (true) and 31 other files (false), expectations by inspection."""
    import base64
    import json
    import os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "_reference_dockerfile_corpus.json")) as f:
        j = json.load(f)
    return [(x["path"], base64.b64decode(x["bytes_b64"]), x["is_dockerfile"]) for x in j["files"]]

# Arguments in JSON form (parseMaybeJSON / parseMaybeJSONToList / parseHealthConfig): a well-formed array with a non-string
# element fails Parse (errDockerfileNotStringArray); whatever json.Decoder.Decode rejects is taken as a plain string.
JSON_KATS = [
    (b'FROM a\nCMD ["x"]\n', True),
    (b'FROM a\nCMD [1]\n', False),                      # errDockerfileNotStringArray
    (b'FROM a\nCMD ["a", 1] tail\n', False),            # Decode reads the first value only
    (b'FROM a\nRUN [1e999]\n', True),                   # number out of range: Decode fails, plain string
    (b'FROM a\nRUN [1,]\n', True),                      # syntax error: plain string
    (b'FROM a\nCOPY [null]\n', False),
    (b'FROM a\nADD [ "a" , "b" ]\n', True),
    (b'FROM a\nHEALTHCHECK CMD [true]\n', False),
    (b'FROM a\nHEALTHCHECK --interval=3s CMD ["x"]\n', True),
    (b'FROM a\nHEALTHCHECK NONE\n', True),
    (b'FROM a\nONBUILD RUN [{}]\n', False),
    (b'FROM a\nRUN []\n', True),
    (b'FROM a\nEXPOSE [1]\n', True),                    # parseStringsWhitespaceDelimited: no JSON
    (b'FROM a\nVOLUME [ "a" , [ ] ]\n', False),
    (b'FROM a\nRUN [1 2]\n', True),
    (b'FROM a\nrun\t[0.5e-3 ]x\n', False),
    (b'FROM a\nRUN ["\\u12g4", 1]\n', True),
    (b'FROM a\nRUN [01]\n', True),
    (b'FROM a\nRUN \\\n  [1]\n', False),                # the array on a continuation line
    (b'FROM a\nRUN [1 \\\n  # c\n ]\n', False),         # ... with a dropped comment inside
    (b'FROM a\nRUN --mount=type=cache [1]\n', False),   # after the flags
    (b'FROM a\nENTRYPOINT [1.7976931348623157e308]\n', False),
    (b'FROM a\nENTRYPOINT [1.7976931348623158e308]\n', False),   # still rounds to MaxFloat64
    (b'FROM a\nENTRYPOINT [1.7976931348623159e308]\n', True),    # rounds to +Inf
    (b'FROM a\nSHELL [{"a":[1e400]}, 2]\n', True),      # the out-of-range number sits two levels down
    (b'FROM a\nSHELL ["a"\x0b]\n', True),               # \v is not JSON white space
    (b'FROM a\nCMD\xc2\xa0[1]\n', True),                # NBSP does not end the command word: the command is unknown
    (b'FROM a\nCMD \xc2\xa0[1]\n', False),              # ... after a blank it is trimmed off the arguments
]

KATS += [("json_%02d" % i, data, want) for i, (data, want) in enumerate(JSON_KATS)]
