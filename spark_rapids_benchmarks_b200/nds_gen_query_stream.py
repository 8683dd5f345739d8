"""`split_special_query` - the one function of the reference's stream generator that sits on the
power-run path (/root/reference/nds/nds_gen_query_stream.py:91-103; called from
nds_power.gen_sql_from_stream :68).  dsqgen itself is absent; see streams.py for the stand-in."""


def split_special_query(q):
    """Split a two-statement template (query14/23/24/39) into its _part1 / _part2 statements.

    `q` is the stream text after '-- start': a header line naming `<template>.tpl`, statement one,
    ';', statement two, ';', and the '-- end ...' trailer.  Part 1 keeps header + first statement with
    the template name rewritten to `<template>_part1.tpl`; part 2 is the rewritten header line followed
    by the second statement.  Both end with ';'."""
    pieces = q.split(';')
    first = pieces[0].replace('.tpl', '_part1.tpl') + ';'
    header = pieces[0].split('\n')[0]
    second = header.replace('.tpl', '_part2.tpl') + '\n' + pieces[1] + ';'
    return first, second
