"""TEST-ONLY backend that answers the facade's device questions from the CPU oracle, so the host
logic of shellac_b200.server.HttpParser can be exercised on a machine without a GPU.  It lives in
tests/ on purpose: the product never constructs it and has no CPU path."""
from shellac_b200.batch import Arena, MessageView
from tests import oracle_lib as O


class OracleBackend(object):
    def batch(self, bufs, frame=False):
        ar = Arena.pack(bufs)
        po = O.orc_parse(ar)
        return po, (O.orc_frame(ar, po) if frame else None)

    def views(self, bufs, frame=False):
        ar = Arena.pack(bufs)
        po = O.orc_parse(ar)
        fo = O.orc_frame(ar, po) if frame else None
        return [MessageView(bufs[i], po, i, fo) for i in range(len(bufs))]

    def chunk_header(self, buf):
        hl, sz, ok = O.orc_chunk_header(Arena.pack([buf]))
        return int(hl[0]), int(sz[0]), bool(ok[0])
