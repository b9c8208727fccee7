"""Regenerate tests/golden/streams_jvm.json: the per-series byte streams of a slice and the records read back from them,
produced by EXECUTING the reference's own bytecode (cramtools-3.0.jar) with oracle/minijvm.py.  TEST INFRASTRUCTURE ONLY.

What runs, unmodified, out of the jar:
  Sam2CramRecordFactory.createCramRecord                          flags, mate fields, read features
  CramSerilization.setMateInfo (cramtools)                        mate linking / detaching            CramSerilization.java:274-373
  CompressionHeaderFactory.build                                  series -> encoding / compressor map [:82-124]
  ContainerFactory.buildSlice                                     AP deltas, Writer.write per record  [:129-205]
    DataWriterFactory.buildWriter (reflection over @DataSeries), Writer.write [:139-261], External*Codec,
    ByteArrayStopCodec, ITF8, ContentDigests (CRC32 / SHA-1 / SHA-512 via zlib / hashlib)
  DataReaderFactory.buildReader, CramRecordReader.read            streams -> records                  [CramRecordReader.java:59-...]
Only ExternalCompressor.compress is replaced: the hook keeps the raw bytes (no gzip in the mini-JVM) and notes which
compressor the header assigned to the series.

    python oracle/make_golden_streams.py        # needs /root/reference; a few minutes
"""
import base64
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

import make_golden_records as G  # noqa: E402
from make_golden_records import vm, jlist, new, boxed, jbytes, b64, FAC, SR, CCR, CHF, features_of  # noqa: E402
from minijvm import JArray, JObject, JavaThrow, i8  # noqa: E402

CF = 'htsjdk/samtools/cram/build/ContainerFactory'
CS = 'htsjdk/samtools/cram/CramSerilization'
CH = 'htsjdk/samtools/cram/structure/CompressionHeader'
RDR = 'htsjdk/samtools/cram/encoding/reader/CramRecordReader'
DRF = 'htsjdk/samtools/cram/encoding/reader/DataReaderFactory'


for _lvl in ('debug', 'info', 'warn', 'error'):          # htsjdk's logger wants SimpleDateFormat and System.err: not part of the path
    vm.hooks[('htsjdk/samtools/util/Log', _lvl, '([Ljava/lang/Object;)V')] = lambda vmx, m, args: None


SFH = 'htsjdk/samtools/SAMFileHeader'
CONTIGS = ['chr1', 'chr2', 'chr3', 'chr4']
READ_GROUPS = ['rgA', 'rgB']


def make_header(ref_len):
    """a real SAMFileHeader: createCramRecord re-resolves the reference indices from the names through it (setHeader, [:109])"""
    h = new(SFH)
    for i, nm in enumerate(CONTIGS):
        sq = new('htsjdk/samtools/SAMSequenceRecord', '(Ljava/lang/String;I)V', nm, ref_len if i == 0 else 10000)
        vm.call(SFH, 'addSequence', '(Lhtsjdk/samtools/SAMSequenceRecord;)V', [h, sq])
    for rg in READ_GROUPS:
        vm.call(SFH, 'addReadGroup', '(Lhtsjdk/samtools/SAMReadGroupRecord;)V',
                [h, new('htsjdk/samtools/SAMReadGroupRecord', '(Ljava/lang/String;)V', rg)])
    return h


def samrecord(rd, header):
    r = G.samrecord(rd)
    vm.call(SR, 'setHeader', '(L%s;)V' % SFH, [r, header])
    vm.call(SR, 'setReferenceName', '(Ljava/lang/String;)V', [r, CONTIGS[rd['ref_id']] if rd['ref_id'] >= 0 else '*'])
    if 'mate_ref' in rd:
        vm.call(SR, 'setMateReferenceName', '(Ljava/lang/String;)V', [r, CONTIGS[rd['mate_ref']] if rd['mate_ref'] >= 0 else '*'])
        vm.call(SR, 'setMateAlignmentStart', '(I)V', [r, rd['mate_pos']])
        vm.call(SR, 'setInferredInsertSize', '(I)V', [r, rd['tlen']])
    if rd.get('rg'):
        vm.call(SR, 'setAttribute', '(Ljava/lang/String;Ljava/lang/Object;)V', [r, 'RG', rd['rg']])
    return r


def add_pairs(ref, reads):
    """proper pairs, a pair whose TLEN is off (detached by setMateInfo), a pair with the mate outside, a secondary, triplets"""
    rng = np.random.default_rng(4242)
    out = []

    def rd(name, flag, pos, mapq, ln, mate_pos, tlen, mate_ref=0):
        b = ref[pos - 1:pos - 1 + ln]
        out.append(dict(name=name, flag=flag, ref_id=0, pos=pos, mapq=mapq, cigar='%dM' % ln, bases=b,
                        quals=bytes(rng.integers(2, 42, size=ln, dtype=np.uint8)), mate_ref=mate_ref, mate_pos=mate_pos, tlen=tlen))

    # proper pairs: first forward (0x63), last reverse (0x93); TLEN = rightmost end - leftmost start + 1
    for i, (p1, p2) in enumerate([(2310, 2500), (2320, 2410), (2400, 2650), (2600, 2601)]):
        L = 60
        t = p2 + L - p1
        rd('pair%d' % i, 0x63, p1, 60, L, p2, t)
        rd('pair%d' % i, 0x93, p2, 60, L, p1, -t)
    rd('badtlen', 0x63, 2700, 50, 50, 2800, 999); rd('badtlen', 0x93, 2800, 50, 50, 2700, -999)          # TLEN disagrees -> detached
    rd('mate_away', 0x63, 2750, 30, 50, 5000, 2300)                                                      # mate not in the container
    rd('mate_unmapped', 0x69, 2760, 30, 50, 2760, 0)                                                     # mate unmapped (0x8)
    rd('sec', 0x163, 2770, 0, 50, 2900, 180); rd('sec', 0x193, 2900, 0, 50, 2770, -180)                  # secondary alignments pair up separately
    rd('trip', 0x63, 2780, 60, 40, 2860, 120); rd('trip', 0x93, 2860, 60, 40, 2780, -120); rd('trip', 0x63, 2950, 60, 40, 2860, 120)
    rd('othercontig', 0x63, 2790, 60, 40, 77, 0, mate_ref=3)
    for i, r in enumerate(out):
        if i % 3 == 1:
            r['rg'] = READ_GROUPS[i % 2]
    return out


def hook_compress(captured):
    def h(vmx, m, args):
        comp, data = args
        captured.append((comp.cls, bytes(data.a)))
        return data
    return h


def jstr(v):
    return v.fields['s'] if isinstance(v, JObject) and v.cls == 'java/lang/String' else v


def get_static(cname, fname):
    vm.load(cname)
    return vm.statics[(cname, fname)]


def list_items(o):
    return o.fields['list']


def enum_name(o):
    return None if o is None else o.fields['name']


def build(reads, ref, coordinate_sorted, tag):
    header0 = make_header(len(ref))
    fac = new(FAC, '([BL%s;Lhtsjdk/samtools/cram/common/Version;)V' % SFH, jbytes(ref), header0, G.make_factory(ref).fields['version'])
    recs = []
    for i, rd in enumerate(reads):
        cr = vm.call(FAC, 'createCramRecord', '(L%s;)L%s;' % (SR, CCR), [fac, samrecord(rd, header0)])
        cr.fields['index'] = i + 1                                       # CramSerilization.java:155
        cr.fields['alignmentStart'] = rd['pos']                          # :156
        qs = cr.fields.get('qualityScores')
        vm.call(CCR, 'setForcePreserveQualityScores', '(Z)V', [cr, 0 if qs is G.NULL_QUALS else 1])     # :160-161 (lossless)
        recs.append(cr)
    lst = jlist(recs)
    if coordinate_sorted:
        vm.call(CS, 'setMateInfo', '(Ljava/util/List;)Ljava/util/List;', [lst])                          # :170-171
    else:
        for cr in recs:
            vm.call(CCR, 'setDetached', '(Z)V', [cr, 1])                                                 # :173-175
    scalars = [dict(flags=cr.fields.get('flags', 0), cf=i8(vm.call(CCR, 'getCompressionFlags', '()B', [cr])) & 0xFF,
                    mate_flags=i8(vm.call(CCR, 'getMateFlags', '()B', [cr])) & 0xFF, read_length=cr.fields.get('readLength', 0),
                    read_group=cr.fields.get('readGroupID', 0), seq_id=cr.fields.get('sequenceId', 0),
                    mate_seq_id=cr.fields.get('mateSequenceID', 0), mate_start=cr.fields.get('mateAlignmentStart', 0),
                    tlen=cr.fields.get('templateSize', 0), nf=cr.fields.get('recordsToNextFragment', 0),
                    mapq=cr.fields.get('mappingQuality', 0), name=jstr(cr.fields.get('readName')))
               for cr in recs]
    header = vm.call(CHF, 'build', '(Ljava/util/List;Lhtsjdk/samtools/cram/structure/SubstitutionMatrix;Z)L%s;' % CH,
                     [new(CHF), lst, None, 1 if coordinate_sorted else 0])
    captured = []
    vm.hooks[('*', 'compress', '([B)[B')] = hook_compress(captured)
    try:
        sl = vm.call(CF, 'buildSlice', '(Ljava/util/List;L%s;)Lhtsjdk/samtools/cram/structure/Slice;' % CH, [lst, header])
    finally:
        del vm.hooks[('*', 'compress', '([B)[B')]
    # the series, by content id
    series = {}
    for k, (key, blk) in sl.fields['external'].fields['map'].items():
        series[int(key.fields['value'])] = bytes(blk.fields['rawContent'].a)
    # which compressor each content id got (R0): class of the anonymous ExternalCompressor + its method / order
    comps = {}
    for k, (key, comp) in header.fields['externalCompressors'].fields['map'].items():
        comps[int(key.fields['value'])] = dict(cls=comp.cls.split('/')[-1], method=enum_name(comp.fields.get('method')),
                                               order=enum_name(comp.fields.get('val$order')))
    enc = {}
    for k, (key, ep) in header.fields['encodingMap'].fields['map'].items():
        enc[key.fields['name']] = dict(id=enum_name(ep.fields['id']), params=b64(ep.fields['params'].a))
    out = dict(tag=tag, coordinate_sorted=bool(coordinate_sorted), reads=[dict(r, bases=b64(r['bases']), quals=b64(r['quals'])) for r in reads],
               scalars=scalars, features=[features_of(cr) for cr in recs],
               series={str(k): b64(v) for k, v in sorted(series.items())}, core=b64(sl.fields['coreBlock'].fields['rawContent'].a),
               compressors={str(k): v for k, v in sorted(comps.items())}, encodings=enc,
               slice=dict(sequence_id=sl.fields.get('sequenceId'), alignment_start=sl.fields.get('alignmentStart'),
                          alignment_span=sl.fields.get('alignmentSpan'), n_records=sl.fields.get('nofRecords'), bases=int(sl.fields.get('bases', 0))),
               ap_delta=int(header.fields.get('APDelta', 0)), read_names_included=int(header.fields.get('readNamesIncluded', 0)), subst_matrix=b64(vm.call(G.SM, 'getEncodedMatrix', '()[B', [header.fields['substitutionMatrix']]).a))
    # slice tags = the content digests (8f-3)
    tags = []
    t = sl.fields.get('sliceTags')
    while t is not None:
        v = t.fields['value']
        tg = t.fields['tag']
        tags.append([chr(tg & 0xFF) + chr((tg >> 8) & 0xFF), b64(v.a) if isinstance(v, JArray) else int(v.fields['value']) if isinstance(v, JObject) else v])
        t = t.fields.get('next')
    out['slice_tags'] = tags
    # slice reference MD5 as CramSerilization.convert sets it after buildContainer (CramSerilization.java:184-197):
    # skipped for a multi-reference slice, MD5 of nothing for an unmapped one, else Slice.setRefMD5(reference) [Slice.java:145-159]
    # -> SequenceUtil.calculateMD5(ref, alignmentStart - 1, min(span, ref.length - alignmentStart + 1)) [SequenceUtil.java:962-966].
    # (SequenceUtil's static initialiser compiles a regex the interpreter's `re` cannot; calculateMD5 does not touch it.)
    vm.skip_clinit.add('htsjdk/samtools/util/SequenceUtil')
    SL = 'htsjdk/samtools/cram/structure/Slice'
    if sl.fields.get('alignmentSpan', 0) < 0:
        sl.fields['alignmentSpan'] = 0
    sid = sl.fields.get('sequenceId')
    if sid == -2:
        out['ref_md5'] = None
    else:
        sl.fields['refMD5'] = None
        try:
            vm.call(SL, 'setRefMD5', '([B)V', [sl, jbytes(b'' if sid == -1 else ref)])
        except NotImplementedError:
            pass                     # the debug log after the assignment [Slice.java:161-175] formats with BigInteger; refMD5 is set by then
        out['ref_md5'] = b64(bytes(sl.fields['refMD5'].a))
    # SURVEY 8f-5 (the block layer around the coded series): Block.write [Block.java:272-294] of the external blocks the header
    # gives to rANS -- method, content type, ITF8 content id / compressed size / raw size, data, CRC32 (little endian) of all of it
    BLK = 'htsjdk/samtools/cram/structure/Block'
    blocks = {}
    for cid, raw in sorted(series.items()):
        cm = comps.get(cid)
        if not cm or cm.get('order') is None or len(raw) == 0 or len(raw) > 6000:
            continue
        order = 1 if cm['order'] == 'ONE' else 0
        comp = vm.call('htsjdk/samtools/cram/io/ExternalCompression', 'rans', '([BI)[B', [jbytes(raw), order])
        blk = new(BLK)
        vm.call(BLK, 'setContentType', '(Lhtsjdk/samtools/cram/structure/BlockContentType;)V',
                [blk, get_static('htsjdk/samtools/cram/structure/BlockContentType', 'EXTERNAL')])
        vm.call(BLK, 'setContentId', '(I)V', [blk, cid])
        vm.call(BLK, 'setMethod', '(Lhtsjdk/samtools/cram/structure/BlockCompressionMethod;)V',
                [blk, get_static('htsjdk/samtools/cram/structure/BlockCompressionMethod', 'RANS')])
        vm.call(BLK, 'setContent', '([B[B)V', [blk, jbytes(raw), comp])
        baos = JObject('java/io/ByteArrayOutputStream')
        vm.jdk_special('java/io/ByteArrayOutputStream', '<init>', '()V', [baos])
        vm.call(BLK, 'write', '(ILjava/io/OutputStream;)V', [blk, 3, baos])
        blocks[str(cid)] = dict(order=order, block=b64(bytes(vm.jdk_virtual('java/io/ByteArrayOutputStream', 'toByteArray', '()[B', [baos]).a)))
    out['blocks_v3'] = blocks

    # SURVEY 8f-5 (the container around the blocks): ContainerIO.writeContainer [ContainerIO.java:152-209] of a container holding
    # this one slice -- container header (ContainerHeaderIO.writeContainerHeader [:69-84]: int32 byte size, ITF8 / LTF8 fields,
    # landmarks, CRC32), compression header block (CompressionHeader.toByteArray), then SliceIO.write [:132-146]: slice header
    # block (createSliceHeaderBlockContent [:76-107], with the content digest tags), core block, external blocks in the map's
    # order.  The external blocks carry what the compress hook returned (the raw bytes) under the method the header chose.
    CT = 'htsjdk/samtools/cram/structure/Container'
    grc = {'all_reads_sorted_header': 0, 'single_ref_sorted': 162, 'unsorted_header': 6000000123}.get(tag, 7)
    cont = new(CT)
    cont.fields['header'] = header
    cont.fields['nofRecords'] = len(recs)
    cont.fields['globalRecordCounter'] = grc
    cont.fields['bases'] = int(sl.fields.get('bases', 0))
    cont.fields['blockCount'] = 0
    cont.fields['slices'] = JArray('L', [sl])
    sl.fields['globalRecordCounter'] = grc
    if cont.fields.get('sequenceId', -1) == -1 and sl.fields.get('sequenceId', 0) != -1:          # ContainerFactory.buildContainer [:93-94]
        cont.fields['sequenceId'] = sl.fields.get('sequenceId', 0)
    vm.call(CF, 'calculateAlignmentBoundaries', '(L%s;)V' % CT, [cont])
    baos = JObject('java/io/ByteArrayOutputStream')
    vm.jdk_special('java/io/ByteArrayOutputStream', '<init>', '()V', [baos])
    vm.call('htsjdk/samtools/cram/structure/ContainerIO', 'writeContainer',
            '(Lhtsjdk/samtools/cram/common/Version;L%s;Ljava/io/OutputStream;)I' % CT, [G.make_factory(ref).fields['version'], cont, baos])
    whole = bytes(vm.jdk_virtual('java/io/ByteArrayOutputStream', 'toByteArray', '()[B', [baos]).a)
    hb = sl.fields['headerBlock']
    out['container_v3'] = dict(bytes=b64(whole), global_record_counter=grc, byte_size=int(cont.fields['containerByteSize']),
                               block_count=int(cont.fields['blockCount']), landmarks=[int(v) for v in cont.fields['landmarks'].a],
                               sequence_id=int(cont.fields.get('sequenceId', 0)), alignment_start=int(cont.fields.get('alignmentStart', 0)),
                               alignment_span=int(cont.fields.get('alignmentSpan', 0)),
                               slice_header_content=b64(bytes(hb.fields['rawContent'].a)),
                               content_ids=[int(v) for v in sl.fields['contentIDs'].a],
                               compression_header=b64(bytes(vm.call(CH, 'toByteArray', '()[B', [header]).a)))

    # ---- read it back: DataReaderFactory.buildReader + CramRecordReader.read (row P)
    imap = JObject('java/util/HashMap'); imap.fields['map'] = {}
    for cid, raw in series.items():
        bis = JObject('java/io/ByteArrayInputStream')
        vm.jdk_special('java/io/ByteArrayInputStream', '<init>', '([B)V', [bis, jbytes(raw)])
        vm.jdk_virtual('java/util/Map', 'put', '', [imap, boxed(cid), bis])
    core = JObject('java/io/ByteArrayInputStream')
    vm.jdk_special('java/io/ByteArrayInputStream', '<init>', '([B)V', [core, jbytes(bytes(sl.fields['coreBlock'].fields['rawContent'].a))])
    bitin = new('htsjdk/samtools/cram/io/DefaultBitInputStream', '(Ljava/io/InputStream;)V', core)
    reader = new(RDR)
    vm.call(DRF, 'buildReader', '(Lhtsjdk/samtools/cram/encoding/reader/AbstractReader;Lhtsjdk/samtools/cram/io/BitInputStream;Ljava/util/Map;L%s;I)Lhtsjdk/samtools/cram/encoding/reader/AbstractReader;' % CH,
            [new(DRF), reader, bitin, imap, header, sl.fields.get('sequenceId')])
    back = []
    read_crs = []
    prev_start = sl.fields.get('alignmentStart')
    for i in range(len(recs)):
        cr = new(CCR)
        vm.call(RDR, 'read', '(L%s;)V' % CCR, [reader, cr])
        if header.fields.get('APDelta', 0):                                 # ContainerParser.getRecords [:196-198]
            prev_start = prev_start + cr.fields.get('alignmentDelta', 0)
            cr.fields['alignmentStart'] = prev_start
        qs = cr.fields.get('qualityScores')
        rb = cr.fields.get('readBases')
        # the CIGAR Cram2SamRecordFactory rebuilds from the read features [Cram2SamRecordFactory.java:119-225]
        cg = vm.call('htsjdk/samtools/cram/build/Cram2SamRecordFactory', 'getCigar2', '(Ljava/util/Collection;I)Lhtsjdk/samtools/Cigar;',
                     [cr.fields.get('readFeatures'), cr.fields.get('readLength', 0)])
        cigar2 = ''.join('%d%s' % (e.fields['length'], 'MIDNSHP=X'[['M', 'I', 'D', 'N', 'S', 'H', 'P', 'EQ', 'X'].index(enum_name(e.fields['operator']))])
                         for e in list_items(cg.fields['cigarElements']))
        # ... and the BAM record cramtools' flat-buffer writer lays out from these values (BAMRecordView.java:70-241, in the order
        # ReaderToBAM.read drives it [ReaderToBAM.java:64-176]); index bin as SAMRecord.computeIndexingBin [SAMRecord.java:1524-1531]
        rd0 = reads[i]; sc0 = scalars[i]
        BV = 'htsjdk/samtools/cram/encoding/reader/BAMRecordView'
        view = new(BV, '([B)V', jbytes(bytes(4096)))
        vm.call(BV, 'setRefID', '(I)V', [view, rd0['ref_id']])
        vm.call(BV, 'setAlignmentStart', '(I)V', [view, rd0['pos']])
        vm.call(BV, 'setReadName', '([B)V', [view, jbytes(rd0['name'].encode('latin-1'))])
        vm.call(BV, 'setMappingScore', '(I)V', [view, rd0['mapq']])
        unm = bool(rd0['flag'] & 4)
        if unm:
            cg_rec = new('htsjdk/samtools/Cigar')
        else:
            cg_rec = cg
        beg = rd0['pos'] - 1
        end = 0 if unm else rd0['pos'] + vm.call('htsjdk/samtools/Cigar', 'getReferenceLength', '()I', [cg_rec]) - 1
        if end <= 0:
            end = beg + 1
        bam_bin = vm.call('htsjdk/samtools/GenomicIndexUtil', 'regionToBin', '(II)I', [beg, end])
        vm.call(BV, 'setIndexBin', '(I)V', [view, bam_bin])
        vm.call(BV, 'setCigar', '(Lhtsjdk/samtools/Cigar;)V', [view, cg_rec])
        vm.call(BV, 'setFlags', '(I)V', [view, rd0['flag']])
        paired = bool(rd0['flag'] & 1)
        vm.call(BV, 'setMateRefID', '(I)V', [view, sc0['mate_seq_id'] if paired else -1])
        vm.call(BV, 'setMateAlStart', '(I)V', [view, sc0['mate_start'] if paired else 0])
        vm.call(BV, 'setInsertSize', '(I)V', [view, sc0['tlen']])
        bb = b'' if rd0['bases'] is None else bytes(rd0['bases'])
        vm.call(BV, 'setBases', '([BII)V', [view, jbytes(bb), 0, len(bb)])
        qq = b'' if rd0['quals'] is None else bytes(rd0['quals'])
        vm.call(BV, 'setQualityScores', '([BII)V', [view, jbytes(qq), 0, len(qq)])
        tagb = (b'RGZ' + rd0['rg'].encode() + b'\0') if rd0.get('rg') else b''
        vm.call(BV, 'setTagData', '([BII)V', [view, jbytes(tagb), 0, len(tagb)])
        bsz = vm.call(BV, 'finish', '()I', [view])
        bam_rec = bytes(view.fields['buf'].a[:bsz])
        read_crs.append(cr)
        back.append(dict(bam=b64(bam_rec), bam_bin=bam_bin, bam_tags=b64(tagb), flags=cr.fields.get('flags', 0), cf=i8(cr.fields.get('compressionFlags', 0)) & 0xFF, seq_id=cr.fields.get('sequenceId', 0),
                         read_length=cr.fields.get('readLength', 0), alignment_delta=cr.fields.get('alignmentDelta', 0),
                         alignment_start=cr.fields.get('alignmentStart', 0),
                         read_group=cr.fields.get('readGroupID', 0), name=jstr(cr.fields.get('readName')),
                         mate_flags=i8(cr.fields.get('mateFlags', 0)) & 0xFF, mate_seq_id=cr.fields.get('mateSequenceID', 0),
                         mate_start=cr.fields.get('mateAlignmentStart', 0), tlen=cr.fields.get('templateSize', 0),
                         nf=cr.fields.get('recordsToNextFragment', 0), mapq=cr.fields.get('mappingQuality', 0),
                         features=features_of(cr), cigar2=cigar2, quals=None if qs is None else b64(qs.a), bases=None if rb is None else b64(rb.a)))
    out['read_back'] = back
    # the pairing part of the decoder (the inverse of setMateInfo): CramNormalizer.normalize [CramNormalizer.java:67-143] on the
    # records as the reader left them -- next / previous links from CF and NF, restoreMateInfo / setNextMate / computeInsertSize
    # [:236-270, 455-470] (it also restores bases and scores: pinned elsewhere)
    norm = new(G.NORM, '(L%s;)V' % SFH, header0)
    lst = JObject('java/util/ArrayList'); lst.fields['list'] = list(read_crs)
    try:
        vm.call(G.NORM, 'normalize', '(Ljava/util/ArrayList;[BIL%s;)V' % G.SM, [norm, lst, jbytes(ref), 0, header.fields['substitutionMatrix']])
    except JavaThrow:
        pass            # a record that overhangs the reference ends the base restoration [:335-434]; the pairing passes ran before it
    idx = {id(c): k for k, c in enumerate(read_crs)}
    out['normalized_mates'] = [dict(flags=c.fields.get('flags', 0), mate_flags=i8(c.fields.get('mateFlags', 0)) & 0xFF, mate_seq_id=c.fields.get('mateSequenceID', 0),
                                    mate_start=c.fields.get('mateAlignmentStart', 0), tlen=c.fields.get('templateSize', 0),
                                    nf=c.fields.get('recordsToNextFragment', 0), alignment_start=c.fields.get('alignmentStart', 0),
                                    next=idx.get(id(c.fields.get('next')), -1), previous=idx.get(id(c.fields.get('previous')), -1))
                               for c in read_crs]
    left = {}
    for k, (key, bis) in imap.fields['map'].items():
        left[str(int(key.fields['value']))] = bis.fields['count'] - bis.fields['pos']
    out['bytes_left'] = left            # the reader must have used up every byte of every series
    return out


def main():
    ref, reads = G.build_reads()
    mapped = [r for r in reads if r['ref_id'] >= 0]
    pairs = add_pairs(ref, reads)
    out = {'generator': 'oracle/make_golden_streams.py (mini-JVM run of cramtools-3.0.jar)', 'ref': b64(ref), 'containers': []}
    # A: the whole golden container plus pairs, input order, coordinate-sorted header (AP deltas incl. negative ones; multi-reference slice: RI)
    out['containers'].append(build(reads + pairs, ref, True, 'all_reads_sorted_header'))
    print('container A done', vm.nbytecodes, flush=True)
    # B: one reference, sorted by position, mates linked
    srt = sorted(mapped + pairs, key=lambda r: r['pos'])
    out['containers'].append(build(srt, ref, True, 'single_ref_sorted'))
    print('container B done', vm.nbytecodes, flush=True)
    # C: unsorted header: absolute AP, every record detached
    out['containers'].append(build(mapped[:40] + pairs, ref, False, 'unsorted_header'))
    print('container C done', vm.nbytecodes, flush=True)
    path = os.path.join(os.path.dirname(HERE), 'tests', 'golden', 'streams_jvm.json')
    with open(path, 'w') as f:
        json.dump(out, f)
    print('wrote', path, os.path.getsize(path), 'bytes; bytecodes executed:', vm.nbytecodes)


if __name__ == '__main__':
    main()
