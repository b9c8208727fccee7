"""Regenerate tests/golden/records_jvm.json by EXECUTING the reference's own bytecode
(cramtools-3.0.jar) with oracle/minijvm.py.  TEST INFRASTRUCTURE ONLY.

What runs, unmodified, out of the jar:
  htsjdk.samtools.cram.build.Sam2CramRecordFactory.createCramRecord(SAMRecord)        (rows F2-F4)
  htsjdk.samtools.cram.build.CompressionHeaderFactory.buildFrequencies / updateSubstitutionCodes,
  htsjdk.samtools.cram.structure.SubstitutionMatrix(long[][]) / getEncodedMatrix      (row F5)
  htsjdk.samtools.cram.lossy.QualityScorePreservation.lossyFromSpec / addQualityScores (row Q; cramtools overlay)
  htsjdk.samtools.cram.build.CramNormalizer.restoreReadBases / restoreQualityScores    (rows N1, N2; cramtools overlay)
  htsjdk.samtools.cram.io.ITF8.writeUnsignedITF8                                       (row W, integer coding)

    python oracle/make_golden_records.py        # needs /root/reference; a minute or two
"""
import base64
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

from minijvm import MiniJVM, JArray, JObject, JavaThrow, REFERENCE_JAR, i8  # noqa: E402
from cramtools_b200 import synth  # noqa: E402

SR = 'htsjdk/samtools/SAMRecord'
CCR = 'htsjdk/samtools/cram/structure/CramCompressionRecord'
FAC = 'htsjdk/samtools/cram/build/Sam2CramRecordFactory'
CHF = 'htsjdk/samtools/cram/build/CompressionHeaderFactory'
SM = 'htsjdk/samtools/cram/structure/SubstitutionMatrix'
QSP = 'htsjdk/samtools/cram/lossy/QualityScorePreservation'
NORM = 'htsjdk/samtools/cram/build/CramNormalizer'
TRACKS = 'htsjdk/samtools/cram/ref/ReferenceTracks'

vm = MiniJVM([REFERENCE_JAR])


def jbytes(b):
    return JArray('B', bytearray(b))


def new(cls, desc='()V', *args):
    vm.load(cls)
    o = JObject(cls)
    vm.call(cls, '<init>', desc, [o] + list(args))
    return o


def jlist(items, cls='java/util/ArrayList'):
    o = JObject(cls)
    o.fields['list'] = list(items)
    return o


def boxed(v):
    return vm.jdk_static_call('java/lang/Integer', 'valueOf', '(I)Ljava/lang/Integer;', [v])


def parse_cigar(s):
    out, num = [], ''
    for ch in s:
        if ch.isdigit():
            num += ch
        else:
            out.append((int(num), ch)); num = ''
    return out


def jcigar(ops):
    vm.load('htsjdk/samtools/CigarOperator')
    els = []
    for ln, op in ops:
        opo = vm.call('htsjdk/samtools/CigarOperator', 'characterToEnum', '(I)Lhtsjdk/samtools/CigarOperator;', [ord(op)])
        els.append(new('htsjdk/samtools/CigarElement', '(ILhtsjdk/samtools/CigarOperator;)V', ln, opo))
    return new('htsjdk/samtools/Cigar', '(Ljava/util/List;)V', jlist(els))


vm.load(SR)
NULL_QUALS = vm.statics[(SR, 'NULL_QUALS')]
NULL_SEQUENCE = vm.statics[(SR, 'NULL_SEQUENCE')]


def samrecord(rd):
    r = new(SR, '(Lhtsjdk/samtools/SAMFileHeader;)V', None)
    vm.call(SR, 'setReadName', '(Ljava/lang/String;)V', [r, rd['name']])
    vm.call(SR, 'setFlags', '(I)V', [r, rd['flag']])
    r.fields['mReferenceIndex'] = boxed(rd['ref_id'])
    r.fields['mMateReferenceIndex'] = boxed(-1)
    vm.call(SR, 'setAlignmentStart', '(I)V', [r, rd['pos']])
    vm.call(SR, 'setMappingQuality', '(I)V', [r, rd['mapq']])
    vm.call(SR, 'setCigar', '(Lhtsjdk/samtools/Cigar;)V', [r, jcigar(parse_cigar(rd['cigar']))])
    vm.call(SR, 'setReadBases', '([B)V', [r, NULL_SEQUENCE if rd['bases'] is None else jbytes(rd['bases'])])
    vm.call(SR, 'setBaseQualities', '([B)V', [r, NULL_QUALS if rd['quals'] is None else jbytes(rd['quals'])])
    return r


def make_factory(ref):
    vm.load(FAC)
    fac = JObject(FAC)   # the constructor wants a SAMFileHeader; only these fields are used on the path
    fac.fields.update(refBases=jbytes(ref), header=None, captureAllTags=0, preserveReadNames=1, refSNPs=None,
                      baseCount=0, featureCount=0, landedRefMaskScores=0, landedTotalScores=0)
    for nm in ('captureTags', 'ignoreTags'):
        st = JObject('java/util/TreeSet'); st.fields['map'] = {}; fac.fields[nm] = st
    fac.fields['readTagList'] = jlist([])
    mp = JObject('java/util/HashMap'); mp.fields['map'] = {}; fac.fields['readGroupMap'] = mp
    vm.load('htsjdk/samtools/cram/common/CramVersions')
    fac.fields['version'] = vm.statics[('htsjdk/samtools/cram/common/CramVersions', 'CRAM_v3')]
    return fac


def feature_json(f):
    t = f.cls.split('/')[-1]
    d = f.fields
    if t == 'Substitution':
        return ['X', d['position'], d.get('base', 0) & 0xFF, d.get('referenceBase', 0) & 0xFF, i8(d.get('code', 0))]
    if t == 'InsertBase':
        return ['i', d['position'], d['base'] & 0xFF]
    if t == 'Deletion':
        return ['D', d['position'], d['length']]
    if t == 'RefSkip':
        return ['N', d['position'], d['length']]
    if t == 'HardClip':
        return ['H', d['position'], d['length']]
    if t == 'Padding':
        return ['P', d['position'], d['length']]
    if t == 'SoftClip':
        return ['S', d['position'], base64.b64encode(bytes(d['sequence'].a)).decode()]
    if t == 'Insertion':
        return ['I', d['position'], base64.b64encode(bytes(d['sequence'].a)).decode()]
    if t == 'BaseQualityScore':
        return ['Q', d['position'], d['qualityScore'] & 0xFF]
    if t == 'ReadBase':
        return ['B', d['position'], d['base'] & 0xFF, d['qualityScore'] & 0xFF]
    raise ValueError(t)


def features_of(cr):
    rf = cr.fields.get('readFeatures')
    return [feature_json(f) for f in rf.fields['list']] if rf is not None else None


def b64(b):
    return None if b is None else base64.b64encode(bytes(b)).decode()


def build_reads():
    rng = np.random.default_rng(77)
    ref = bytearray(synth.reference(4000, 31).tobytes())
    ref[500:560] = bytes(ref[500:560]).lower()          # soft-masked stretch: normalizeBase upper-cases
    ref[900:930] = b'N' * 30                            # assembly gap
    ref[1200:1203] = b'RYK'                             # IUPAC codes -> N
    ref = bytes(ref)
    reads = []

    def add(name, flag, pos, mapq, cigar, bases, quals, ref_id=0):
        reads.append(dict(name=name, flag=flag, ref_id=ref_id, pos=pos, mapq=mapq, cigar=cigar, bases=bases, quals=quals))

    def q(n, lo=2, hi=41):
        return bytes(rng.integers(lo, hi + 1, size=n, dtype=np.uint8))

    def mut(b, rate=0.03):
        b = bytearray(b)
        for i in range(len(b)):
            if rng.random() < rate:
                b[i] = b'ACGTN'[rng.integers(0, 5)]
        return bytes(b)

    # hand-made shapes
    add('perfect', 0, 101, 60, '50M', ref[100:150], q(50))
    add('subs', 16, 201, 60, '60M', mut(ref[200:260], 0.1), q(60))
    add('lowercase_ref', 0, 491, 30, '80M', ref[490:570].upper(), q(80))
    add('lowercase_ref_read_lc', 0, 491, 3, '80M', ref[490:570], q(80))
    add('ngap', 0, 881, 60, '70M', mut(ref[880:950].replace(b'N', b'A')), q(70))
    add('iupac', 0, 1191, 4, '30M', ref[1190:1220], q(30))
    add('ins', 0, 301, 60, '20M3I27M', ref[300:320] + b'GGA' + ref[320:347], q(50))
    add('del', 0, 401, 60, '25M2D25M', ref[400:425] + ref[427:452], q(50))
    add('del_at_end_flank', 0, 601, 60, '49M5D1M', ref[600:649] + ref[654:655], q(50))
    add('softclips', 0, 701, 60, '7S36M7S', b'TTTTTTT' + mut(ref[700:736]) + b'GGGGGGG', q(50))
    add('hardclip', 0, 751, 60, '5H40M3H', ref[750:790], q(40))
    add('refskip', 0, 1001, 60, '20M100N20M', ref[1000:1020] + ref[1120:1140], q(40))
    add('padding', 0, 1301, 60, '10M2P10M', ref[1300:1320], q(20))
    add('eq_x_ops', 0, 1401, 60, '10=1X9=', ref[1400:1410] + b'N' + ref[1411:1420], q(20))
    add('complex', 16, 1501, 37, '4S10M1I5M2D10M3I8M5S', b'ACGT' + ref[1500:1510] + b'C' + ref[1510:1515] + ref[1517:1527]
        + b'TTT' + ref[1527:1535] + b'NNNNN', q(46))
    add('overhang_end', 0, 3981, 60, '40M', ref[3980:4000] + b'A' * 20, q(40))
    add('star_quals', 0, 1601, 60, '30M', mut(ref[1600:1630]), None)
    add('no_bases', 0, 1701, 60, '30M', None, None)
    add('low_mapq', 0, 1801, 2, '50M', mut(ref[1800:1850], 0.05), q(50))
    add('high_quals', 0, 1901, 60, '50M', mut(ref[1900:1950]), q(50, 38, 93))
    add('mostly_mismatch', 0, 2001, 60, '30M', mut(ref[2000:2030], 0.9), q(30))
    add('ins_first', 0, 2101, 60, '2I28M', b'AA' + ref[2100:2128], q(30))
    add('unmapped_flag_with_pos', 4, 2201, 0, '30M', ref[2200:2230], q(30))
    add('unmapped', 4, 0, 0, '', bytes(rng.choice(np.frombuffer(b'ACGTN', np.uint8), size=40)), q(40), ref_id=-1)
    # simulated reads with many events
    sim = synth.reads(np.frombuffer(ref, np.uint8), 120, 99, read_len=100, sub_rate=0.02, indel_frac=0.3, clip_frac=0.2,
                      contig_start=2300, contig_end=3950)
    for i in range(120):
        lo, hi = sim['read_off'][i], sim['read_off'][i + 1]
        cg = ''.join('%d%s' % c for c in synth.cigar_of(sim, i))
        add('sim%d' % i, int(sim['flag'][i]), int(sim['pos'][i]), int(sim['mapq'][i]), cg,
            sim['bases'][lo:hi].tobytes(), sim['quals'][lo:hi].tobytes())
    return ref, reads


SPECS = ['*8', 'm5', 'M30', '*40', '*0', 'N40-D8', 'R8-N40', 'I40', 'N8', 'D40-I8', 'm20_8-N40', 'U']


def main():
    ref, reads = build_reads()
    out = {'generator': 'oracle/make_golden_records.py (mini-JVM run of cramtools-3.0.jar)',
           'ref': b64(ref), 'reads': [dict(r, bases=b64(r['bases']), quals=b64(r['quals'])) for r in reads]}
    fac = make_factory(ref)
    recs = []
    for rd in reads:
        cr = vm.call(FAC, 'createCramRecord', '(L%s;)L%s;' % (SR, CCR), [fac, samrecord(rd)])
        cr.fields['alignmentStart'] = rd['pos']      # CramSerilization.java:156
        recs.append(cr)
    out['features'] = [features_of(cr) for cr in recs]
    out['unknown_bases'] = [int(vm.call(CCR, 'isUnknownBases', '()Z', [cr])) for cr in recs]
    out['bf_flags'] = [cr.fields['flags'] for cr in recs]

    # substitution matrix (row F5)
    freqs = vm.call(CHF, 'buildFrequencies', '(Ljava/util/List;)[[J', [jlist(recs)])
    sm = new(SM, '([[J)V', freqs)
    out['subst_matrix'] = b64(vm.call(SM, 'getEncodedMatrix', '()[B', [sm]).a)
    out['subst_counts'] = {'%c%c' % (r, b): int(freqs.a[r].a[b]) for r in b'ACGTN' for b in b'ACGTN' if freqs.a[r].a[b]}
    vm.call(CHF, 'updateSubstitutionCodes', '(Ljava/util/List;L%s;)V' % SM, [jlist(recs), sm])
    out['features_coded'] = [features_of(cr) for cr in recs]
    # extra matrices: ties, zero counts, single substitutions
    extra = []
    rng = np.random.default_rng(5)
    for t in range(12):
        counts = {}
        vm.load(SM)
        fr = vm.multi_array('[[J', [256, 256])
        for r in b'ACGTN':
            for b in b'ACGTN':
                if r != b:
                    c = int(rng.choice([0, 0, 1, 1, 2, 5, 100, 100, 7]))
                    fr.a[r].a[b] = c
                    if c:
                        counts['%c%c' % (r, b)] = c
        m = new(SM, '([[J)V', fr)
        extra.append({'counts': counts, 'matrix': b64(vm.call(SM, 'getEncodedMatrix', '()[B', [m]).a)})
    out['subst_extra'] = extra

    # lossy quality models (row Q)
    tracks = new(TRACKS, '(ILjava/lang/String;[B)V', 0, 'chr1', jbytes(ref))
    lossy = {}
    for spec in SPECS:
        qsp = vm.call(QSP, 'lossyFromSpec', '(Ljava/lang/String;)L%s;' % QSP, [spec])
        rows = []
        for rd in reads:
            s = samrecord(rd)
            cr = vm.call(FAC, 'createCramRecord', '(L%s;)L%s;' % (SR, CCR), [fac, s])
            cr.fields['alignmentStart'] = rd['pos']
            try:
                vm.call(QSP, 'addQualityScores', '(L%s;L%s;L%s;)V' % (SR, CCR, TRACKS), [qsp, s, cr, tracks])
                qs = cr.fields.get('qualityScores')
                rows.append({'force': int(vm.call(CCR, 'isForcePreserveQualityScores', '()Z', [cr])),
                             'scores': None if qs is None else ('*' if qs is NULL_QUALS else b64(qs.a)),
                             'features': features_of(cr)})
            except JavaThrow as t:
                rows.append({'throws': '%s' % t.obj.cls})
        lossy[spec] = rows
        print('lossy', spec, 'done', vm.nbytecodes, flush=True)
    out['lossy'] = lossy

    # decode side (rows N1, N2): region = [min start, max end] of the mapped records, as restoreBases builds it
    mapped = [i for i, rd in enumerate(reads) if rd['ref_id'] >= 0 and rd['pos'] > 0]
    for name, ws, we in (('full_span', 100, 4000), ('cut_short', 100, 3990), ('late_start', 101, 4000)):
        window = ref[ws:we]
        matrix = sm
        rows = []
        for i, rd in enumerate(reads):
            # a record as the reader leaves it: features with codes, no bases
            s = samrecord(rd)
            cr = vm.call(FAC, 'createCramRecord', '(L%s;)L%s;' % (SR, CCR), [fac, s])
            cr.fields['alignmentStart'] = rd['pos']
            vm.call(CHF, 'updateSubstitutionCodes', '(Ljava/util/List;L%s;)V' % SM, [jlist([cr]), sm])
            rf = cr.fields.get('readFeatures')
            if rf is not None:
                for f in rf.fields['list']:
                    if f.cls.endswith('Substitution'):
                        f.fields['base'] = 0; f.fields['referenceBase'] = 0
            if rd['ref_id'] < 0 or rd['pos'] == 0 or (rd['flag'] & 4):
                rows.append(None)
                continue
            try:
                b = vm.call(NORM, 'restoreReadBases', '(L%s;[BIL%s;)[B' % (CCR, SM), [cr, jbytes(window), ws, matrix])
                rows.append(b64(b.a))
            except JavaThrow as t:
                rows.append({'throws': t.obj.cls})
        out.setdefault('restore_bases', {})[name] = {'window_start0': ws, 'window_end0': we, 'bases': rows}
        print('restore', name, 'done', flush=True)

    # quality restoration after each lossy model (+ lossless)
    rq = {}
    for spec in ['lossless'] + SPECS:
        qsp = None if spec == 'lossless' else vm.call(QSP, 'lossyFromSpec', '(Ljava/lang/String;)L%s;' % QSP, [spec])
        rows = []
        for rd in reads:
            s = samrecord(rd)
            cr = vm.call(FAC, 'createCramRecord', '(L%s;)L%s;' % (SR, CCR), [fac, s])
            try:
                if qsp is None:
                    vm.call(CCR, 'setForcePreserveQualityScores', '(Z)V', [cr, 0 if cr.fields['qualityScores'] is NULL_QUALS else 1])
                else:
                    vm.call(QSP, 'addQualityScores', '(L%s;L%s;L%s;)V' % (SR, CCR, TRACKS), [qsp, s, cr, tracks])
                qs = cr.fields.get('qualityScores')
                if qs is not None and qs is not NULL_QUALS:
                    cr.fields['qualityScores'] = JArray('B', bytearray(qs.a))     # what the reader would hand over
                res = vm.call(NORM, 'restoreQualityScores', '(BL%s;)[B' % CCR, [ord('?') - ord('!'), cr])
                rows.append('*' if res is NULL_QUALS else b64(res.a))
            except JavaThrow as t:
                rows.append({'throws': t.obj.cls})
        rq[spec] = rows
    out['restore_quals'] = rq

    # ITF8 (row W)
    itf = []
    for v in [0, 1, 127, 128, 16383, 16384, 2097151, 2097152, 268435455, 268435456, 2147483647, -1, -2, -129, -2147483648,
              150, 10000, 99, 163, 147, 83]:
        bb = vm.jdk_static_call('java/nio/ByteBuffer', 'allocate', '(I)Ljava/nio/ByteBuffer;', [8])
        vm.call('htsjdk/samtools/cram/io/ITF8', 'writeUnsignedITF8', '(ILjava/nio/ByteBuffer;)V', [v, bb])
        itf.append([v, b64(bb.arr[:bb.position])])
    out['itf8'] = itf

    path = os.path.join(os.path.dirname(HERE), 'tests', 'golden', 'records_jvm.json')
    with open(path, 'w') as f:
        json.dump(out, f)
    print('wrote', path, os.path.getsize(path), 'bytes; bytecodes executed:', vm.nbytecodes)


if __name__ == '__main__':
    main()
