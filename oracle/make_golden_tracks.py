"""Regenerate tests/golden/tracks_jvm.json: the lossy selectors X (lower than coverage) and P (pile-up of mismatches) of
SURVEY 8f-4, produced by EXECUTING the reference's own bytecode (cramtools-3.0.jar) with oracle/minijvm.py.
TEST INFRASTRUCTURE ONLY.

What runs, unmodified, out of the jar, container after container on ONE ReferenceTracks object, as CramSerilization.convert
does (src/main/java/htsjdk/samtools/cram/CramSerilization.java:112-166):
  CramSerilization.getSpans                 the container's alignment span per sequence            :60-86
  ReferenceTracks.ensureRange               window of the coverage / mismatch tracks                ref/ReferenceTracks.java:133-139
  CramSerilization.updateTracks             coverage and mismatch counts of the container's records :225-257
  Sam2CramRecordFactory.createCramRecord, QualityScorePreservation.addQualityScores (with the tracks)
The ReferenceTracks are built with the class's private window-size constructor so that a 4 kbp reference exercises window
moves, the clamp at the reference end and the cut-off beyond the window.

    python oracle/make_golden_tracks.py        # needs /root/reference; a few minutes
"""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

import make_golden_records as G  # noqa: E402
from make_golden_records import vm, jlist, new, jbytes, b64, FAC, SR, CCR, TRACKS, QSP, NULL_QUALS  # noqa: E402
from minijvm import JavaThrow  # noqa: E402

CS = 'htsjdk/samtools/cram/CramSerilization'
SPECS = ['X3', 'X8_8', 'P0_40', 'P1_8', 'N40-X5', 'P0_40-X4_8-D8', 'R8-X2', 'X6P1_40']
WINDOWS = [1000000, 1500, 700]
PER_CONTAINER = 12

for _lvl in ('debug', 'info', 'warn', 'error'):
    vm.hooks[('htsjdk/samtools/util/Log', _lvl, '([Ljava/lang/Object;)V')] = lambda vmx, m, args: None
vm.hooks[('java/io/PrintStream', 'println', '(Ljava/lang/String;)V')] = lambda vmx, m, args: None     # "moving tracks to ..."


def scenario(ref, conts, window, spec, fac, want_tracks):
    tracks = new(TRACKS, '(ILjava/lang/String;[BI)V', 0, 'chr1', jbytes(ref), window)
    qsp = vm.call(QSP, 'lossyFromSpec', '(Ljava/lang/String;)L%s;' % QSP, [spec])
    out = []
    for reads in conts:
        sams = [G.samrecord(rd) for rd in reads]
        row = {}
        try:
            spans = vm.call(CS, 'getSpans', '(Ljava/util/List;)Ljava/util/Map;', [jlist(sams)])
            span = [v for k, (key, v) in spans.fields['map'].items() if key.fields['value'] == 0][0]
            row['span'] = [int(span.fields.get('start', 0)), int(span.fields.get('span', 0))]
            vm.call(TRACKS, 'ensureRange', '(II)V', [tracks, row['span'][0], row['span'][1]])
            row['position_after_ensure'] = int(tracks.fields['position'])
            vm.call(CS, 'updateTracks', '(Ljava/util/List;L%s;)V' % TRACKS, [jlist(sams), tracks])
            if want_tracks:
                row['coverage'] = [int(v) for v in tracks.fields['coverage'].a]
                row['mismatches'] = [int(v) for v in tracks.fields['mismatches'].a]
            recs = []
            for rd, s in zip(reads, sams):
                cr = vm.call(FAC, 'createCramRecord', '(L%s;)L%s;' % (SR, CCR), [fac, s])
                cr.fields['alignmentStart'] = rd['pos']
                try:
                    vm.call(QSP, 'addQualityScores', '(L%s;L%s;L%s;)V' % (SR, CCR, TRACKS), [qsp, s, cr, tracks])
                    qs = cr.fields.get('qualityScores')
                    recs.append({'force': int(vm.call(CCR, 'isForcePreserveQualityScores', '()Z', [cr])),
                                 'scores': None if qs is None else ('*' if qs is NULL_QUALS else b64(qs.a)),
                                 'position': int(tracks.fields['position'])})
                except JavaThrow as t:
                    recs.append({'throws': '%s' % t.obj.cls, 'position': int(tracks.fields['position'])})
            row['records'] = recs
            row['position_at_end'] = int(tracks.fields['position'])
        except JavaThrow as t:
            row['throws'] = '%s' % t.obj.cls
            out.append(row)
            break
        out.append(row)
    return out


def main():
    ref, reads = G.build_reads()
    mapped = sorted([r for r in reads if r['ref_id'] >= 0 and r['bases'] is not None], key=lambda r: r['pos'])
    conts = [mapped[i:i + PER_CONTAINER] for i in range(0, len(mapped), PER_CONTAINER)]
    fac = G.make_factory(ref)
    out = {'generator': 'oracle/make_golden_tracks.py (mini-JVM run of cramtools-3.0.jar)', 'ref': b64(ref),
           'containers': [[dict(r, bases=b64(r['bases']), quals=None if r['quals'] is None else b64(r['quals'])) for r in c] for c in conts],
           'scenarios': []}
    for w in WINDOWS:
        for k, spec in enumerate(SPECS):
            out['scenarios'].append({'window': w, 'spec': spec, 'containers': scenario(ref, conts, w, spec, fac, k == 0)})
            print('window', w, spec, 'done', vm.nbytecodes, flush=True)
    path = os.path.join(os.path.dirname(HERE), 'tests', 'golden', 'tracks_jvm.json')
    with open(path, 'w') as f:
        json.dump(out, f)
    print('wrote', path, os.path.getsize(path), 'bytes; bytecodes executed:', vm.nbytecodes)


if __name__ == '__main__':
    main()
