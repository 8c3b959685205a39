"""Import shim for boltons.strutils (only names, never on the numeric path)."""
import re


def slugify(text, delim="_", lower=True, ascii=False):
    out = re.sub(r"[^\w]+", delim, str(text)).strip(delim)
    return out.lower() if lower else out


def a10n(string):
    return string if len(string) < 3 else f"{string[0]}{len(string) - 2}{string[-1]}"


def asciify(text, ignore=False):
    return str(text).encode("ascii", "ignore" if ignore else "replace")


def format_int_list(int_list, delim=",", range_delim="-", delim_space=False):
    return (delim + (" " if delim_space else "")).join(str(i) for i in int_list)


def iter_splitlines(text):
    yield from text.splitlines()


def parse_int_list(range_string, delim=",", range_delim="-"):
    out = []
    for part in range_string.split(delim):
        if range_delim in part:
            a, b = part.split(range_delim)
            out.extend(range(int(a), int(b) + 1))
        elif part.strip():
            out.append(int(part))
    return sorted(out)


def unwrap_text(text, ending="\n\n"):
    return text
