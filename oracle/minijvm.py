"""oracle/minijvm.py -- TEST INFRASTRUCTURE ONLY.

A deliberately small Java bytecode interpreter, just large enough to EXECUTE the
reference's own rANS classes (htsjdk/samtools/cram/encoding/rans/*.class and a few
other leaf classes) straight out of /root/reference/cramtools-3.0.jar.  There is no
JVM in the build container, so this is how golden vectors for the hot path are
produced from the reference itself rather than from a restatement:

    python oracle/make_golden.py      # runs the jar's bytecode, writes tests/golden/*.json

It is slow (pure Python, ~1 M bytecodes/s), so vectors are small.  Nothing at run time
(tests on the GPU box, smoke(), bench.py) imports this module or needs the jar: the
vectors are committed.

Supported: the class-file format (constant pool, fields, methods, Code attribute,
exception tables), ints/longs/doubles/floats with Java wrap-around semantics, arrays, objects,
static init, enums, tableswitch/lookupswitch, exceptions thrown by `athrow` (surfaced
as JavaThrow), and emulations of the handful of JDK classes the rANS code touches
(java.nio.ByteBuffer/ByteOrder, System.arraycopy, Arrays.fill, StringBuilder, Math).
"""
import struct
import zipfile

# ----------------------------------------------------------------------------- class file


class ClassFile:
    def __init__(self, data):
        self.data = data
        self.pos = 0
        magic = self.u4()
        assert magic == 0xCAFEBABE
        self.u2(); self.u2()
        n = self.u2()
        self.cp = [None] * n
        i = 1
        while i < n:
            tag = self.u1()
            if tag == 1:
                ln = self.u2()
                self.cp[i] = ('utf8', self.data[self.pos:self.pos + ln].decode('utf-8', 'replace'))
                self.pos += ln
            elif tag == 3:
                self.cp[i] = ('int', struct.unpack('>i', self.raw(4))[0])
            elif tag == 4:
                self.cp[i] = ('float', struct.unpack('>f', self.raw(4))[0])
            elif tag == 5:
                self.cp[i] = ('long', struct.unpack('>q', self.raw(8))[0]); i += 1
            elif tag == 6:
                self.cp[i] = ('double', struct.unpack('>d', self.raw(8))[0]); i += 1
            elif tag == 7:
                self.cp[i] = ('class', self.u2())
            elif tag == 8:
                self.cp[i] = ('string', self.u2())
            elif tag in (9, 10, 11):
                self.cp[i] = ({9: 'field', 10: 'method', 11: 'imethod'}[tag], self.u2(), self.u2())
            elif tag == 12:
                self.cp[i] = ('nat', self.u2(), self.u2())
            elif tag == 15:
                self.cp[i] = ('mh', self.u1(), self.u2())
            elif tag == 16:
                self.cp[i] = ('mt', self.u2())
            elif tag == 18:
                self.cp[i] = ('indy', self.u2(), self.u2())
            else:
                raise ValueError('cp tag %d' % tag)
            i += 1
        self.access = self.u2()
        self.name = self.class_name(self.u2())
        sup = self.u2()
        self.super_name = self.class_name(sup) if sup else None
        self.interfaces = [self.class_name(self.u2()) for _ in range(self.u2())]
        self.fields = []
        for _ in range(self.u2()):
            acc, nm, ds = self.u2(), self.utf(self.u2()), self.utf(self.u2())
            attrs = self.attrs()
            self.fields.append((acc, nm, ds, attrs))
        self.methods = {}
        for _ in range(self.u2()):
            acc, nm, ds = self.u2(), self.utf(self.u2()), self.utf(self.u2())
            attrs = self.attrs()
            self.methods[(nm, ds)] = Method(self, acc, nm, ds, attrs)

    def raw(self, n):
        b = self.data[self.pos:self.pos + n]; self.pos += n; return b

    def u1(self):
        v = self.data[self.pos]; self.pos += 1; return v

    def u2(self):
        v = struct.unpack_from('>H', self.data, self.pos)[0]; self.pos += 2; return v

    def u4(self):
        v = struct.unpack_from('>I', self.data, self.pos)[0]; self.pos += 4; return v

    def utf(self, i):
        return self.cp[i][1]

    def class_name(self, i):
        return self.utf(self.cp[i][1])

    def attrs(self):
        out = {}
        for _ in range(self.u2()):
            nm = self.utf(self.u2()); ln = self.u4()
            out[nm] = self.raw(ln)
        return out

    def member_ref(self, i):
        _, ci, nti = self.cp[i]
        _, ni, di = self.cp[nti]
        return self.class_name(ci), self.utf(ni), self.utf(di)


class Method:
    def __init__(self, cf, acc, name, desc, attrs):
        self.cf, self.acc, self.name, self.desc = cf, acc, name, desc
        self.static = bool(acc & 0x0008)
        self.code = None
        self.lines = []
        if 'Code' in attrs:
            a = attrs['Code']
            self.max_stack, self.max_locals, cl = struct.unpack_from('>HHI', a, 0)
            self.code = a[8:8 + cl]
            p = 8 + cl
            n = struct.unpack_from('>H', a, p)[0]; p += 2
            self.extable = []
            for _ in range(n):
                self.extable.append(struct.unpack_from('>HHHH', a, p)); p += 8
            na = struct.unpack_from('>H', a, p)[0]; p += 2
            for _ in range(na):
                ni, ln = struct.unpack_from('>HI', a, p); p += 6
                if cf.utf(ni) == 'LineNumberTable':
                    k = struct.unpack_from('>H', a, p)[0]
                    for t in range(k):
                        self.lines.append(struct.unpack_from('>HH', a, p + 2 + 4 * t))
                p += ln
        self.nargs = self._count_args(desc)

    @staticmethod
    def _count_args(desc):
        """list of slot widths of the parameters"""
        out = []; i = 1
        while desc[i] != ')':
            c = desc[i]
            if c in 'JD':
                out.append(2); i += 1
            elif c == 'L':
                out.append(1); i = desc.index(';', i) + 1
            elif c == '[':
                while desc[i] == '[':
                    i += 1
                if desc[i] == 'L':
                    i = desc.index(';', i) + 1
                else:
                    i += 1
                out.append(1)
            else:
                out.append(1); i += 1
        return out

    def line_of(self, pc):
        best = None
        for spc, ln in self.lines:
            if spc <= pc:
                best = ln
        return best


# ----------------------------------------------------------------------------- runtime objects


class JavaThrow(Exception):
    def __init__(self, obj):
        self.obj = obj
        Exception.__init__(self, '%s: %s' % (obj.cls, obj.fields.get('message')))


class JObject:
    __slots__ = ('cls', 'fields')

    def __init__(self, cls):
        self.cls = cls
        self.fields = {}

    def __repr__(self):
        return '<%s %r>' % (self.cls, self.fields)


class JArray:
    __slots__ = ('kind', 'a')

    def __init__(self, kind, a):
        self.kind, self.a = kind, a


def i32(v):
    v &= 0xFFFFFFFF
    return v - 0x100000000 if v & 0x80000000 else v


def i64(v):
    v &= 0xFFFFFFFFFFFFFFFF
    return v - 0x10000000000000000 if v & 0x8000000000000000 else v


def i8(v):
    v &= 0xFF
    return v - 256 if v & 0x80 else v


def i16(v):
    v &= 0xFFFF
    return v - 65536 if v & 0x8000 else v


def d2i(d, bits):
    if d != d:
        return 0
    lo, hi = -(1 << (bits - 1)), (1 << (bits - 1)) - 1
    if d >= hi:
        return hi
    if d <= lo:
        return lo
    return int(d)


class ByteOrder:
    def __init__(self, name):
        self.name = name


BIG = ByteOrder('BIG_ENDIAN')
LITTLE = ByteOrder('LITTLE_ENDIAN')


class ByteBuffer:
    """java.nio.HeapByteBuffer semantics, the subset the CRAM code uses."""

    def __init__(self, arr, offset, cap):
        self.arr = arr          # python bytearray shared with JArray('B')
        self.offset = offset
        self.cap = cap
        self.position = 0
        self.limit = cap
        self.order = BIG
        self.jarray = None

    def _chk(self, vm, n):
        if self.limit - self.position < n:
            vm.throw('java/nio/BufferUnderflowException' , None)

    def _chkw(self, vm, n):
        if self.limit - self.position < n:
            vm.throw('java/nio/BufferOverflowException', None)


# ----------------------------------------------------------------------------- the VM


class MiniJVM:
    def __init__(self, jar_paths):
        self.jars = [zipfile.ZipFile(p) for p in jar_paths]
        self.classes = {}
        self.statics = {}
        self.initialized = set()
        self.nbytecodes = 0
        self.strings = {}
        self.skip_clinit = set()      # classes whose static initialiser is not run (it needs JDK features the interpreter lacks and the test does not touch)
        self.hooks = {}      # (class, method, descriptor) or ('*', method, descriptor) -> python function(vm, method, args): replaces a jar method

    # ---- loading
    def load(self, name):
        cf = self.classes.get(name)
        if cf is None:
            data = None
            for z in self.jars:
                try:
                    data = z.read(name + '.class'); break
                except KeyError:
                    pass
            if data is None:
                raise KeyError('class not in jar: ' + name)
            cf = ClassFile(data)
            self.classes[name] = cf
        if name not in self.initialized:
            self.initialized.add(name)
            if cf.super_name and not cf.super_name.startswith('java/'):
                self.load(cf.super_name)
            for acc, nm, ds, attrs in cf.fields:
                if acc & 0x0008:
                    v = 0.0 if ds in 'DF' else (0 if ds in 'IJSBCZ' else None)
                    if 'ConstantValue' in attrs:
                        ci = struct.unpack('>H', attrs['ConstantValue'])[0]
                        v = cf.cp[ci][1]
                    self.statics[(name, nm)] = v
            m = cf.methods.get(('<clinit>', '()V'))
            if m and name not in self.skip_clinit:
                self.invoke(m, [])
        return cf

    def is_jdk(self, name):
        return name.startswith('java/')

    def jdk_base(self, cname):
        """first JDK class among the superclasses of a jar class (ExposedByteArrayOutputStream -> ByteArrayOutputStream)"""
        c = cname
        while c and not self.is_jdk(c):
            c = self.load(c).super_name
        return c

    # ---- just enough java.lang.reflect for DataWriterFactory.buildWriter / DataReaderFactory.buildReader
    def public_fields(self, cname):
        """Class.getFields(): public fields of the class and its superclasses, as java.lang.reflect.Field stand-ins that
        carry the RuntimeVisibleAnnotations of the class file (enum-valued and string-valued elements)"""
        out = []
        c = cname
        while c and not self.is_jdk(c):
            cf = self.load(c)
            for acc, nm, ds, attrs in cf.fields:
                if not acc & 0x0001:
                    continue
                f = JObject('java/lang/reflect/Field')
                f.fields.update(owner=c, name=nm, desc=ds, static=bool(acc & 0x0008), annotations=self._annotations(cf, attrs))
                out.append(f)
            c = cf.super_name
        return out

    def _annotations(self, cf, attrs):
        raw = attrs.get('RuntimeVisibleAnnotations')
        res = {}
        if not raw:
            return res
        p = 2
        for _ in range(struct.unpack_from('>H', raw, 0)[0]):
            tname = cf.utf(struct.unpack_from('>H', raw, p)[0])[1:-1]
            npairs = struct.unpack_from('>H', raw, p + 2)[0]
            p += 4
            proxy = JObject('java/lang/annotation/$Proxy')
            for _ in range(npairs):
                ename = cf.utf(struct.unpack_from('>H', raw, p)[0])
                tag = chr(raw[p + 2])
                if tag == 'e':
                    ecls = cf.utf(struct.unpack_from('>H', raw, p + 3)[0])[1:-1]
                    econst = cf.utf(struct.unpack_from('>H', raw, p + 5)[0])
                    self.load(ecls)
                    proxy.fields[ename] = self.statics[(ecls, econst)]
                    p += 7
                elif tag == 's':
                    proxy.fields[ename] = cf.utf(struct.unpack_from('>H', raw, p + 3)[0])
                    p += 5
                else:
                    raise NotImplementedError('annotation element tag ' + tag)
            res[tname] = proxy
        return res

    def find_method(self, cname, name, desc):
        c = cname
        while c and not self.is_jdk(c):
            cf = self.load(c)
            m = cf.methods.get((name, desc))
            if m:
                return m
            c = cf.super_name
        return None

    def find_static_owner(self, cname, fname):
        c = cname
        while c and not self.is_jdk(c):
            if (c, fname) in self.statics:
                return c
            cf = self.load(c)
            if (c, fname) in self.statics:
                return c
            c = cf.super_name
        raise KeyError((cname, fname))

    def instance_of(self, obj, cname):
        if obj is None:
            return False
        if isinstance(obj, str):
            return cname in ('java/lang/String', 'java/lang/Object', 'java/lang/CharSequence', 'java/lang/Comparable')
        c = obj.cls if isinstance(obj, JObject) else None
        if isinstance(obj, ByteBuffer):
            return cname in ('java/nio/ByteBuffer', 'java/nio/Buffer', 'java/lang/Object')
        if isinstance(obj, JArray):
            if cname in ('java/lang/Object', 'java/lang/Cloneable', 'java/io/Serializable'):
                return True
            if cname.startswith('[') and len(cname) == 2:                     # primitive array class: [B [S [I ...
                return obj.kind == cname[1] or (obj.kind == 'Z' and cname[1] == 'B' and False)
            return cname.startswith('[') and obj.kind in ('L', '[')         # object arrays: element types are not tracked
        while c:
            if c == cname:
                return True
            if self.is_jdk(c):
                return JDK_SUPERS.get(c) is not None and self.instance_of_jdk(c, cname)
            cf = self.load(c)
            if cname in cf.interfaces:
                return True
            c = cf.super_name
        return False

    def instance_of_jdk(self, c, cname):
        while c:
            if c == cname:
                return True
            c = JDK_SUPERS.get(c)
        return False

    # ---- helpers
    def new_string(self, s):
        return s  # java.lang.String modelled as python str

    def throw(self, cname, msg):
        o = JObject(cname)
        o.fields['message'] = msg
        raise JavaThrow(o)

    def new_array(self, atype, n):
        if n < 0:
            self.throw('java/lang/NegativeArraySizeException', None)
        if atype in ('B',):
            return JArray('B', bytearray(n))
        if atype in ('Z',):
            return JArray('Z', [0] * n)
        if atype in ('D', 'F'):
            return JArray(atype, [0.0] * n)
        if atype in ('I', 'J', 'S', 'C'):
            return JArray(atype, [0] * n)
        return JArray('L', [None] * n)

    def multi_array(self, desc, dims):
        # desc like [[I
        if len(dims) == 1:
            return self.new_array(desc[1] if desc[1] != 'L' and desc[1] != '[' else 'L', dims[0])
        arr = JArray('L', [None] * dims[0])
        for i in range(dims[0]):
            arr.a[i] = self.multi_array(desc[1:], dims[1:])
        return arr

    # ---- interpreter
    def call(self, cname, name, desc, args):
        self.load(cname)
        m = self.find_method(cname, name, desc)
        if m is None:
            raise KeyError((cname, name, desc))
        return self.invoke(m, args)

    def invoke(self, m, args):
        if self.hooks:
            h = self.hooks.get((m.cf.name, m.name, m.desc)) or self.hooks.get(('*', m.name, m.desc))
            if h is not None:
                return h(self, m, args)
        if m.code is None:
            raise NotImplementedError('native/abstract %s.%s%s' % (m.cf.name, m.name, m.desc))
        loc = [None] * m.max_locals
        i = 0
        if not m.static:
            loc[0] = args[0]; i = 1
            args = args[1:]
        for w, a in zip(m.nargs, args):
            loc[i] = a; i += w
        try:
            return self.run(m, loc)
        except NotImplementedError as e:       # missing JDK emulation: say where the jar needed it
            e.args = (str(e.args[0]) + '\n    at %s.%s%s' % (m.cf.name, m.name, m.desc),) + e.args[1:]
            raise

    def run(self, m, loc):
        code = m.code
        cf = m.cf
        cp = cf.cp
        st = []
        push = st.append
        pop = st.pop
        pc = 0
        up = struct.unpack_from
        while True:
            try:
                while True:
                    op = code[pc]
                    self.nbytecodes += 1
                    # --- constants
                    if op == 0:
                        pc += 1
                    elif op == 1:
                        push(None); pc += 1
                    elif 2 <= op <= 8:
                        push(op - 3); pc += 1
                    elif op in (9, 10):
                        push(op - 9); pc += 1
                    elif 11 <= op <= 13:
                        push(float(op - 11)); pc += 1
                    elif op in (14, 15):
                        push(float(op - 14)); pc += 1
                    elif op == 16:
                        push(i8(code[pc + 1])); pc += 2
                    elif op == 17:
                        push(up('>h', code, pc + 1)[0]); pc += 3
                    elif op in (18, 19, 20):
                        if op == 18:
                            idx = code[pc + 1]; pc += 2
                        else:
                            idx = up('>H', code, pc + 1)[0]; pc += 3
                        e = cp[idx]
                        if e[0] == 'string':
                            push(cf.utf(e[1]))
                        elif e[0] == 'class':
                            push(('class', cf.utf(e[1])))
                        else:
                            push(e[1])
                    # --- loads
                    elif 21 <= op <= 25:
                        push(loc[code[pc + 1]]); pc += 2
                    elif 26 <= op <= 45:
                        push(loc[(op - 26) & 3]); pc += 1
                    elif 46 <= op <= 53:   # array loads
                        idx = pop(); arr = pop()
                        if arr is None:
                            self.throw('java/lang/NullPointerException', None)
                        if idx < 0 or idx >= len(arr.a):
                            self.throw('java/lang/ArrayIndexOutOfBoundsException', str(idx))
                        v = arr.a[idx]
                        if op == 51 and arr.kind == 'B':
                            v = v - 256 if v & 0x80 else v
                        push(v); pc += 1
                    # --- stores
                    elif 54 <= op <= 58:
                        loc[code[pc + 1]] = pop(); pc += 2
                    elif 59 <= op <= 78:
                        loc[(op - 59) & 3] = pop(); pc += 1
                    elif 79 <= op <= 86:   # array stores
                        v = pop(); idx = pop(); arr = pop()
                        if arr is None:
                            self.throw('java/lang/NullPointerException', None)
                        if idx < 0 or idx >= len(arr.a):
                            self.throw('java/lang/ArrayIndexOutOfBoundsException', str(idx))
                        if op == 84:
                            v = v & 0xFF if arr.kind == 'B' else (1 if v & 1 else 0)
                        elif op == 85:
                            v &= 0xFFFF
                        elif op == 86:
                            v = i16(v)
                        arr.a[idx] = v; pc += 1
                    # --- stack
                    elif op == 87:
                        pop(); pc += 1
                    elif op == 88:
                        v = pop()
                        if not self._is_wide(m, v):
                            pop()
                        pc += 1
                    elif op == 89:
                        push(st[-1]); pc += 1
                    elif op == 90:
                        v1 = pop(); v2 = pop(); push(v1); push(v2); push(v1); pc += 1
                    elif op == 91:  # dup_x2 (treat wide values as single python items)
                        v1 = pop(); v2 = pop()
                        if isinstance(v2, WIDE):
                            push(v1); push(v2); push(v1)
                        else:
                            v3 = pop(); push(v1); push(v3); push(v2); push(v1)
                        pc += 1
                    elif op == 92:  # dup2
                        v1 = st[-1]
                        if isinstance(v1, WIDE):
                            push(v1)
                        else:
                            v2 = st[-2]; push(v2); push(v1)
                        pc += 1
                    elif op == 93:  # dup2_x1
                        v1 = pop()
                        if isinstance(v1, WIDE):
                            v2 = pop(); push(v1); push(v2); push(v1)
                        else:
                            v2 = pop(); v3 = pop(); push(v2); push(v1); push(v3); push(v2); push(v1)
                        pc += 1
                    elif op == 95:
                        v1 = pop(); v2 = pop(); push(v1); push(v2); pc += 1
                    # --- int arithmetic
                    elif op == 96:
                        b = pop(); a = pop(); push(i32(a + b)); pc += 1
                    elif op == 97:
                        b = pop(); a = pop(); push(L(i64(a + b))); pc += 1
                    elif op in (98, 99):
                        b = pop(); a = pop(); push(D(a + b) if op == 99 else a + b); pc += 1
                    elif op == 100:
                        b = pop(); a = pop(); push(i32(a - b)); pc += 1
                    elif op == 101:
                        b = pop(); a = pop(); push(L(i64(a - b))); pc += 1
                    elif op in (102, 103):
                        b = pop(); a = pop(); push(D(a - b) if op == 103 else a - b); pc += 1
                    elif op == 104:
                        b = pop(); a = pop(); push(i32(a * b)); pc += 1
                    elif op == 105:
                        b = pop(); a = pop(); push(L(i64(a * b))); pc += 1
                    elif op in (106, 107):
                        b = pop(); a = pop(); push(D(a * b) if op == 107 else a * b); pc += 1
                    elif op in (108, 109):
                        b = pop(); a = pop()
                        if b == 0:
                            self.throw('java/lang/ArithmeticException', '/ by zero')
                        q = abs(a) // abs(b)
                        if (a < 0) != (b < 0):
                            q = -q
                        push(i32(q) if op == 108 else L(i64(q))); pc += 1
                    elif op in (110, 111):
                        b = pop(); a = pop()
                        if b == 0.0:
                            r = float('nan') if a == 0.0 or a != a else (float('inf') if (a > 0) == (str(b)[0] != '-') else float('-inf'))
                        else:
                            r = a / b
                        push(D(r) if op == 111 else r); pc += 1
                    elif op in (112, 113):
                        b = pop(); a = pop()
                        if b == 0:
                            self.throw('java/lang/ArithmeticException', '/ by zero')
                        r = abs(a) % abs(b)
                        if a < 0:
                            r = -r
                        push(i32(r) if op == 112 else L(i64(r))); pc += 1
                    elif op == 116:
                        push(i32(-pop())); pc += 1
                    elif op == 117:
                        push(L(i64(-pop()))); pc += 1
                    elif op in (118, 119):
                        v = -pop(); push(D(v) if op == 119 else v); pc += 1
                    elif op == 120:
                        b = pop(); a = pop(); push(i32(a << (b & 31))); pc += 1
                    elif op == 121:
                        b = pop(); a = pop(); push(L(i64(a << (b & 63)))); pc += 1
                    elif op == 122:
                        b = pop(); a = pop(); push(a >> (b & 31)); pc += 1
                    elif op == 123:
                        b = pop(); a = pop(); push(L(a >> (b & 63))); pc += 1
                    elif op == 124:
                        b = pop(); a = pop(); push(i32((a & 0xFFFFFFFF) >> (b & 31))); pc += 1
                    elif op == 125:
                        b = pop(); a = pop(); push(L(i64((a & 0xFFFFFFFFFFFFFFFF) >> (b & 63)))); pc += 1
                    elif op == 126:
                        b = pop(); a = pop(); push(a & b); pc += 1
                    elif op == 127:
                        b = pop(); a = pop(); push(L(a & b)); pc += 1
                    elif op == 128:
                        b = pop(); a = pop(); push(a | b); pc += 1
                    elif op == 129:
                        b = pop(); a = pop(); push(L(a | b)); pc += 1
                    elif op == 130:
                        b = pop(); a = pop(); push(a ^ b); pc += 1
                    elif op == 131:
                        b = pop(); a = pop(); push(L(a ^ b)); pc += 1
                    elif op == 132:
                        idx = code[pc + 1]; loc[idx] = i32(loc[idx] + i8(code[pc + 2])); pc += 3
                    # --- conversions
                    elif op == 133:
                        push(L(pop())); pc += 1
                    elif op == 134:
                        push(struct.unpack('f', struct.pack('f', float(pop())))[0]); pc += 1
                    elif op == 135:
                        push(D(float(pop()))); pc += 1
                    elif op == 136:
                        push(i32(pop())); pc += 1
                    elif op == 137:
                        push(struct.unpack('f', struct.pack('f', float(pop())))[0]); pc += 1
                    elif op == 138:
                        push(D(float(pop()))); pc += 1
                    elif op == 139:
                        push(d2i(pop(), 32)); pc += 1
                    elif op == 140:
                        push(L(d2i(pop(), 64))); pc += 1
                    elif op == 141:
                        push(D(float(pop()))); pc += 1
                    elif op == 142:
                        push(d2i(pop(), 32)); pc += 1
                    elif op == 143:
                        push(L(d2i(pop(), 64))); pc += 1
                    elif op == 144:
                        push(struct.unpack('f', struct.pack('f', float(pop())))[0]); pc += 1
                    elif op == 145:
                        push(i8(pop())); pc += 1
                    elif op == 146:
                        push(pop() & 0xFFFF); pc += 1
                    elif op == 147:
                        push(i16(pop())); pc += 1
                    # --- compares
                    elif op == 148:
                        b = pop(); a = pop(); push((a > b) - (a < b)); pc += 1
                    elif 149 <= op <= 152:
                        b = pop(); a = pop()
                        if a != a or b != b:
                            push(1 if op in (150, 152) else -1)
                        else:
                            push((a > b) - (a < b))
                        pc += 1
                    elif 153 <= op <= 158:
                        v = pop()
                        t = (v == 0, v != 0, v < 0, v >= 0, v > 0, v <= 0)[op - 153]
                        pc = pc + up('>h', code, pc + 1)[0] if t else pc + 3
                    elif 159 <= op <= 164:
                        b = pop(); a = pop()
                        t = (a == b, a != b, a < b, a >= b, a > b, a <= b)[op - 159]
                        pc = pc + up('>h', code, pc + 1)[0] if t else pc + 3
                    elif op in (165, 166):
                        b = pop(); a = pop()
                        t = (a is b) if op == 165 else (a is not b)
                        if isinstance(a, str) and isinstance(b, str):
                            t = (a == b) if op == 165 else (a != b)
                        pc = pc + up('>h', code, pc + 1)[0] if t else pc + 3
                    elif op == 167:
                        pc += up('>h', code, pc + 1)[0]
                    elif op == 170:   # tableswitch
                        base = pc
                        p = (pc + 4) & ~3
                        dflt, lo, hi = up('>iii', code, p)
                        v = pop()
                        if lo <= v <= hi:
                            pc = base + up('>i', code, p + 12 + 4 * (v - lo))[0]
                        else:
                            pc = base + dflt
                    elif op == 171:   # lookupswitch
                        base = pc
                        p = (pc + 4) & ~3
                        dflt, n = up('>ii', code, p)
                        v = pop()
                        tgt = dflt
                        for k in range(n):
                            mv, off = up('>ii', code, p + 8 + 8 * k)
                            if mv == v:
                                tgt = off; break
                        pc = base + tgt
                    elif 172 <= op <= 176:
                        return pop()
                    elif op == 177:
                        return None
                    # --- fields
                    elif op == 178:
                        cn, fn, ds = cf.member_ref(up('>H', code, pc + 1)[0])
                        if self.is_jdk(cn):
                            push(self.jdk_static(cn, fn))
                        else:
                            self.load(cn)
                            v = self.statics[(self.find_static_owner(cn, fn), fn)]
                            push(self._wrap(ds, v))
                        pc += 3
                    elif op == 179:
                        cn, fn, ds = cf.member_ref(up('>H', code, pc + 1)[0])
                        self.load(cn)
                        self.statics[(self.find_static_owner(cn, fn), fn)] = pop()
                        pc += 3
                    elif op == 180:
                        cn, fn, ds = cf.member_ref(up('>H', code, pc + 1)[0])
                        o = pop()
                        if o is None:
                            self.throw('java/lang/NullPointerException', None)
                        v = o.fields.get(fn, _DEFAULT)
                        if v is _DEFAULT:
                            v = 0.0 if ds in 'DF' else (0 if ds in 'IJSBCZ' else None)
                        push(self._wrap(ds, v)); pc += 3
                    elif op == 181:
                        cn, fn, ds = cf.member_ref(up('>H', code, pc + 1)[0])
                        v = pop(); o = pop()
                        if o is None:
                            self.throw('java/lang/NullPointerException', None)
                        o.fields[fn] = v; pc += 3
                    # --- invokes
                    elif op in (182, 183, 184, 185):
                        cn, mn, ds = cf.member_ref(up('>H', code, pc + 1)[0])
                        widths = Method._count_args(ds)
                        nargs = len(widths) + (0 if op == 184 else 1)
                        args = st[len(st) - nargs:] if nargs else []
                        if nargs:
                            del st[len(st) - nargs:]
                        pc += 5 if op == 185 else 3
                        ret = self.dispatch(op, cn, mn, ds, args)
                        if not ds.endswith(')V'):
                            push(self._wrap(ds[ds.index(')') + 1:], ret))
                    elif op == 187:
                        cn = cf.class_name(up('>H', code, pc + 1)[0])
                        if not self.is_jdk(cn):
                            self.load(cn)
                        push(JObject(cn)); pc += 3
                    elif op == 188:
                        t = {4: 'Z', 5: 'C', 6: 'F', 7: 'D', 8: 'B', 9: 'S', 10: 'I', 11: 'J'}[code[pc + 1]]
                        push(self.new_array(t, pop())); pc += 2
                    elif op == 189:
                        push(self.new_array('L', pop())); pc += 3
                    elif op == 190:
                        a = pop()
                        if a is None:
                            self.throw('java/lang/NullPointerException', None)
                        push(len(a.a)); pc += 1
                    elif op == 191:
                        o = pop()
                        if o is None:
                            self.throw('java/lang/NullPointerException', None)
                        raise JavaThrow(o)
                    elif op == 192:   # checkcast: trust
                        pc += 3
                    elif op == 193:
                        cn = cf.class_name(up('>H', code, pc + 1)[0])
                        push(1 if self.instance_of(pop(), cn) else 0); pc += 3
                    elif op in (194, 195):
                        pop(); pc += 1
                    elif op == 196:   # wide
                        op2 = code[pc + 1]; idx = up('>H', code, pc + 2)[0]
                        if op2 == 132:
                            loc[idx] = i32(loc[idx] + up('>h', code, pc + 4)[0]); pc += 6
                        elif 21 <= op2 <= 25:
                            push(loc[idx]); pc += 4
                        else:
                            loc[idx] = pop(); pc += 4
                    elif op == 197:
                        desc = cf.class_name(up('>H', code, pc + 1)[0]); nd = code[pc + 3]
                        dims = st[len(st) - nd:]; del st[len(st) - nd:]
                        push(self.multi_array(desc, dims)); pc += 4
                    elif op == 198:
                        pc = pc + up('>h', code, pc + 1)[0] if pop() is None else pc + 3
                    elif op == 199:
                        pc = pc + up('>h', code, pc + 1)[0] if pop() is not None else pc + 3
                    elif op == 200:
                        pc += up('>i', code, pc + 1)[0]
                    else:
                        raise NotImplementedError('opcode %d in %s.%s' % (op, cf.name, m.name))
            except JavaThrow as t:
                handler = None
                for s, e, h, ct in m.extable:
                    if s <= pc < e:
                        if ct == 0 or self.instance_of(t.obj, cf.class_name(ct)):
                            handler = h; break
                if handler is None:
                    raise
                del st[:]
                push(t.obj)
                pc = handler

    @staticmethod
    def _is_wide(m, v):
        return isinstance(v, WIDE)

    @staticmethod
    def _wrap(ds, v):
        if ds == 'J' and not isinstance(v, L):
            return L(v)
        if ds == 'D' and not isinstance(v, D):
            return D(v)
        return v

    # ---- method dispatch incl. JDK emulation
    def dispatch(self, op, cn, mn, ds, args):
        if not (mn == '<init>' and cn == 'java/lang/String'):
            # `new String(...)` objects behave as python strings everywhere they are used
            args = [a.fields['s'] if isinstance(a, JObject) and a.cls == 'java/lang/String' else a for a in args]
        if op == 184:
            if self.is_jdk(cn):
                return self.jdk_static_call(cn, mn, ds, args)
            self.load(cn)
            m = self.find_method(cn, mn, ds)
            return self.invoke(m, args)
        recv = args[0]
        if recv is None:
            self.throw('java/lang/NullPointerException', None)
        if op == 183:  # invokespecial: exact class
            if self.is_jdk(cn):
                return self.jdk_special(cn, mn, ds, args)
            m = self.find_method(cn, mn, ds)
            return self.invoke(m, args)
        # virtual / interface
        if isinstance(recv, JObject) and not self.is_jdk(recv.cls):
            m = self.find_method(recv.cls, mn, ds)
            if m is not None:
                return self.invoke(m, args)
        return self.jdk_virtual(cn, mn, ds, args)

    def empty_list(self):
        # Collections.EMPTY_LIST / emptyList(): one shared immutable instance, compared by identity in the reference
        if not hasattr(self, '_empty_list'):
            self._empty_list = JObject('java/util/ArrayList')
            self._empty_list.fields['list'] = []
        return self._empty_list

    def jdk_static(self, cn, fn):
        if cn == 'java/nio/ByteOrder':
            return LITTLE if fn == 'LITTLE_ENDIAN' else BIG
        if cn == 'java/util/Collections' and fn == 'EMPTY_LIST':
            return self.empty_list()
        if cn == 'java/lang/System' and fn in ('out', 'err'):
            return JObject('java/io/PrintStream')
        raise NotImplementedError('static field %s.%s' % (cn, fn))

    def jdk_special(self, cn, mn, ds, args):
        o = args[0]
        if mn == '<init>':
            if cn in ('java/lang/Object', 'java/io/OutputStream', 'java/io/InputStream'):
                return None
            if cn in ('java/io/DataInputStream', 'java/io/FilterInputStream'):
                o.fields['in'] = args[1]
                return None
            if cn == 'java/io/FilterOutputStream':
                o.fields['out'] = args[1]
                return None
            if cn == 'java/io/ByteArrayOutputStream':
                o.fields['buf'] = JArray('B', bytearray()); o.fields['count'] = 0
                return None
            if cn == 'java/io/ByteArrayInputStream':
                o.fields['buf'] = args[1]; o.fields['pos'] = 0; o.fields['count'] = len(args[1].a)
                return None
            if cn == 'java/util/zip/CRC32':
                o.fields['crc'] = 0
                return None
            if cn == 'java/lang/Enum':
                o.fields['name'] = args[1]; o.fields['ordinal'] = args[2]; return None
            if cn == 'java/lang/StringBuilder':
                o.fields['s'] = args[1] if len(args) > 1 and isinstance(args[1], str) else ''
                return None
            if cn == 'java/lang/String':
                a = args[1:]
                if not a:
                    o.fields['s'] = ''
                elif isinstance(a[0], JArray) and a[0].kind == 'C':
                    chars = a[0].a if len(a) == 1 else a[0].a[a[1]:a[1] + a[2]]
                    o.fields['s'] = ''.join(chr(c) for c in chars)
                elif isinstance(a[0], JArray) and a[0].kind == 'B':
                    o.fields['s'] = bytes(a[0].a).decode('latin-1')
                else:
                    o.fields['s'] = a[0] if isinstance(a[0], str) else a[0].fields['s']
                return None
            if cn in JDK_LISTS:
                o.fields['list'] = list(args[1].fields['list']) if len(args) > 1 and isinstance(args[1], JObject) and 'list' in args[1].fields else []
                return None
            if cn in JDK_MAPS:
                o.fields['map'] = {}
                if cn == 'java/util/TreeMap' and ds == '(Ljava/util/Comparator;)V':
                    o.fields['cmp'] = args[1]
                return None
            if cn in JDK_SETS:
                o.fields['map'] = {}
                return None
            if cn in ('java/lang/StringBuffer',):
                o.fields['s'] = args[1] if len(args) > 1 and isinstance(args[1], str) else ''
                return None
            if cn in JDK_SUPERS and self.instance_of_jdk(cn, 'java/lang/Throwable'):
                o.fields['message'] = args[1] if len(args) > 1 and isinstance(args[1], str) else None
                return None
        raise NotImplementedError('special %s.%s%s' % (cn, mn, ds))

    def jdk_static_call(self, cn, mn, ds, args):
        if cn == 'java/lang/System' and mn in ('nanoTime', 'currentTimeMillis'):
            return 0                                    # timings of the jar's own logging: never part of a result
        if cn == 'java/lang/System' and mn == 'arraycopy':
            src, sp, dst, dp, n = args
            if src is None or dst is None:
                self.throw('java/lang/NullPointerException', None)
            if sp < 0 or dp < 0 or n < 0 or sp + n > len(src.a) or dp + n > len(dst.a):
                self.throw('java/lang/ArrayIndexOutOfBoundsException', None)
            dst.a[dp:dp + n] = src.a[sp:sp + n]
            return None
        if cn == 'java/util/Arrays' and mn == 'fill':
            if len(args) == 4:
                a, lo, hi, v = args
                if lo > hi:
                    self.throw('java/lang/IllegalArgumentException', 'fromIndex > toIndex')
                if lo < 0 or hi > len(a.a):
                    self.throw('java/lang/ArrayIndexOutOfBoundsException', None)
            else:
                a, v = args; lo, hi = 0, len(a.a)
            if a.kind == 'B':
                v &= 0xFF
                a.a[lo:hi] = bytes([v]) * (hi - lo)
            else:
                a.a[lo:hi] = [v] * (hi - lo)
            return None
        if cn == 'java/util/Arrays' and mn in ('copyOf', 'copyOfRange'):
            if mn == 'copyOf':
                a, n = args; lo = 0; hi = n
            else:
                a, lo, hi = args
            out = self.new_array(a.kind, hi - lo)
            k = max(0, min(hi, len(a.a)) - lo)
            out.a[0:k] = a.a[lo:lo + k]
            return out
        if cn == 'java/nio/ByteBuffer':
            if mn == 'allocate':
                arr = self.new_array('B', args[0])
                bb = ByteBuffer(arr.a, 0, args[0]); bb.jarray = arr
                return bb
            if mn == 'wrap':
                arr = args[0]
                if len(args) == 1:
                    bb = ByteBuffer(arr.a, 0, len(arr.a))
                else:
                    bb = ByteBuffer(arr.a, 0, len(arr.a)); bb.position = args[1]; bb.limit = args[1] + args[2]
                bb.jarray = arr
                return bb
        if cn == 'java/lang/Math':
            if mn == 'min':
                return min(args)
            if mn == 'max':
                return max(args)
            if mn == 'abs':
                return abs(args[0])
        if cn == 'java/lang/String' and mn == 'format':
            return args[0]            # only log messages are formatted on the paths run here: the pattern stands in for the text
        if cn == 'java/util/Collections':
            if mn == 'emptyList':
                return self.empty_list()
            if mn == 'singletonList':
                o = JObject('java/util/ArrayList'); o.fields['list'] = [args[0]]; return o
            if mn == 'sort':
                lst = args[0].fields['list']
                if len(args) == 2 and args[1] is not None:
                    import functools
                    cmpo = args[1]
                    m = self.find_method(cmpo.cls, 'compare', '(Ljava/lang/Object;Ljava/lang/Object;)I')
                    lst.sort(key=functools.cmp_to_key(lambda a, b: self.invoke(m, [cmpo, a, b])))   # stable, like Java
                else:
                    lst.sort(key=lambda v: v.fields['value'] if isinstance(v, JObject) else v)
                return None
            if mn in ('unmodifiableList', 'unmodifiableMap', 'unmodifiableSet', 'unmodifiableCollection'):
                return args[0]
        if cn in ('java/lang/Character', 'java/lang/Byte', 'java/lang/Short', 'java/lang/Long', 'java/lang/Boolean') and mn == 'valueOf':
            o = JObject(cn); o.fields['value'] = args[0]; return o
        if cn == 'java/lang/Character' and mn == 'isDigit':
            return 1 if 48 <= args[0] <= 57 else 0
        if cn == 'java/util/Arrays' and mn == 'sort':
            import functools
            arr = args[0]
            lo, hi = (args[1], args[2]) if len(args) == 4 else (0, len(arr.a))
            cmpo = args[-1] if isinstance(args[-1], JObject) else None
            part = list(arr.a[lo:hi])
            if cmpo is not None:
                m = self.find_method(cmpo.cls, 'compare', '(Ljava/lang/Object;Ljava/lang/Object;)I')
                part.sort(key=functools.cmp_to_key(lambda a, b: self.invoke(m, [cmpo, a, b])))   # stable (TimSort)
            else:
                part.sort()
            arr.a[lo:hi] = part
            return None
        if cn == 'java/util/Arrays' and mn == 'asList':
            o = JObject('java/util/ArrayList'); o.fields['list'] = list(args[0].a); return o
        if cn == 'java/util/Arrays' and mn == 'equals':
            a, b = args
            return 1 if (a is None and b is None) or (a is not None and b is not None and list(a.a) == list(b.a)) else 0
        if cn == 'java/lang/System' and mn == 'getProperty':
            return args[1] if len(args) > 1 else None   # no -D properties set: defaults apply
        if cn == 'java/lang/Integer' and mn == 'parseInt':
            return i32(int(args[0]))
        if cn == 'java/lang/Integer' and mn == 'valueOf':
            o = JObject('java/lang/Integer'); o.fields['value'] = args[0]; return o
        if cn == 'java/lang/String' and mn == 'valueOf':
            return self.to_jstring(args[0], ds)
        if cn == 'java/util/regex/Pattern' and mn == 'compile':
            import re as _re
            o = JObject('java/util/regex/Pattern'); o.fields['re'] = _re.compile(args[0]); o.fields['pattern'] = args[0]; return o
        if cn == 'java/nio/charset/Charset' and mn == 'forName':
            o = JObject('java/nio/charset/Charset'); o.fields['name'] = args[0]; return o
        if cn == 'java/util/EnumSet' and mn in ('allOf', 'noneOf', 'of', 'copyOf'):
            o = JObject('java/util/EnumSet'); o.fields['map'] = {}
            if mn == 'allOf':
                self.load(args[0][1])
                vals = self.statics[(args[0][1], '$VALUES')].a
            elif mn == 'noneOf':
                vals = []
            elif mn == 'copyOf':
                vals = list(args[0].fields['map'].values()) if 'map' in args[0].fields else list(args[0].fields['list'])
            else:
                vals = [a for a in args if isinstance(a, JObject)] + [x for a in args if isinstance(a, JArray) for x in a.a]
            for v in sorted(vals, key=lambda e: e.fields['ordinal']):
                o.fields['map'][_key(v)] = v
            return o
        if cn == 'java/security/MessageDigest' and mn == 'getInstance':
            import hashlib
            alg = {'SHA-1': 'sha1', 'SHA-512': 'sha512', 'MD5': 'md5', 'SHA-256': 'sha256'}[args[0]]
            o = JObject('java/security/MessageDigest'); o.fields['alg'] = alg; o.fields['h'] = hashlib.new(alg); return o
        if cn == 'java/lang/Enum' and mn == 'valueOf':
            _, cname = args[0]
            for v in self.statics[(cname, '$VALUES')].a:
                if v.fields['name'] == args[1]:
                    return v
            self.throw('java/lang/IllegalArgumentException', args[1])
        raise NotImplementedError('static %s.%s%s' % (cn, mn, ds))

    def to_jstring(self, v, ds):
        p = ds[1:ds.index(')')]
        if p == 'C':
            return chr(v)
        if p == 'Z':
            return 'true' if v else 'false'
        if v is None:
            return 'null'
        if isinstance(v, JObject):
            if 'name' in v.fields and 'ordinal' in v.fields:
                return v.fields['name']
            return str(v)
        if isinstance(v, float):
            return repr(float(v))
        return str(int(v)) if not isinstance(v, str) else v

    def jdk_virtual(self, cn, mn, ds, args):
        o = args[0]
        if isinstance(o, ByteBuffer):
            return self.bb_call(o, mn, ds, args[1:])
        if isinstance(o, tuple) and o[0] == 'class':
            if mn == 'desiredAssertionStatus':
                return 0       # plain `java` without -ea
            if mn == 'isArray':
                return 1 if o[1].startswith('[') else 0
            if mn == 'getName':
                return o[1].replace('/', '.')
            if mn == 'getSimpleName':
                return o[1].split('/')[-1].split('$')[-1]
        if isinstance(o, tuple) and o[0] == 'class' and mn == 'getFields':
            return JArray('L', self.public_fields(o[1]))
        if isinstance(o, JArray) and mn == 'clone':
            return JArray(o.kind, o.a[:])
        if isinstance(o, JObject):
            base = o.cls if self.is_jdk(o.cls) else self.jdk_base(o.cls)
            if base == 'java/io/ByteArrayOutputStream':
                buf = o.fields['buf'].a
                if mn == 'write':
                    if ds == '(I)V':
                        buf.append(args[1] & 0xFF)
                    elif ds == '([B)V':
                        buf.extend(args[1].a)
                    else:
                        src, sp, n = args[1:]
                        if sp < 0 or n < 0 or sp + n > len(src.a):
                            self.throw('java/lang/IndexOutOfBoundsException', None)
                        buf.extend(src.a[sp:sp + n])
                    o.fields['count'] = len(buf)
                    return None
                if mn == 'toByteArray':
                    return JArray('B', bytearray(buf))
                if mn == 'size':
                    return len(buf)
                if mn == 'reset':
                    del buf[:]; o.fields['count'] = 0; return None
                if mn in ('flush', 'close'):
                    return None
            if base == 'java/io/ByteArrayInputStream':
                buf = o.fields['buf'].a
                if mn == 'read' and ds == '()I':
                    if o.fields['pos'] >= o.fields['count']:
                        return -1
                    v = buf[o.fields['pos']]; o.fields['pos'] += 1; return v
                if mn == 'read':
                    dst, dp, n = (args[1], 0, len(args[1].a)) if ds == '([B)I' else args[1:]
                    left = o.fields['count'] - o.fields['pos']
                    if left <= 0:
                        return -1 if n > 0 else 0
                    n = min(n, left)
                    dst.a[dp:dp + n] = buf[o.fields['pos']:o.fields['pos'] + n]
                    o.fields['pos'] += n
                    return n
                if mn == 'available':
                    return o.fields['count'] - o.fields['pos']
                if mn == 'close':
                    return None
            if base in ('java/io/OutputStream', 'java/io/InputStream') and mn in ('flush', 'close'):
                return None
            if base == 'java/io/OutputStream' and mn == 'write' and ds == '([B)V':
                # OutputStream.write(byte[]) = write(b, 0, b.length), which loops over the subclass's write(int)
                m1 = self.find_method(o.cls, 'write', '([BII)V')
                if m1 is not None:
                    return self.invoke(m1, [o, args[1], 0, len(args[1].a)])
                m1 = self.find_method(o.cls, 'write', '(I)V')
                for b in args[1].a:
                    self.invoke(m1, [o, b])
                return None
            if base == 'java/io/OutputStream' and mn == 'write' and ds == '([BII)V':
                m1 = self.find_method(o.cls, 'write', '(I)V')
                for b in args[1].a[args[2]:args[2] + args[3]]:
                    self.invoke(m1, [o, b])
                return None
            if o.cls == 'java/util/regex/Pattern' and mn == 'matcher':
                r = JObject('java/util/regex/Matcher'); r.fields['re'] = o.fields['re']; r.fields['s'] = args[1]; return r
            if o.cls == 'java/util/regex/Matcher':
                if mn == 'matches':
                    return 1 if o.fields['re'].fullmatch(o.fields['s']) else 0
                if mn == 'find':
                    return 1 if o.fields['re'].search(o.fields['s']) else 0
            if o.cls == 'java/lang/reflect/Field':
                if mn == 'isAnnotationPresent':
                    return 1 if args[1][1] in o.fields['annotations'] else 0
                if mn == 'getAnnotation':
                    return o.fields['annotations'].get(args[1][1])
                if mn == 'set':
                    args[1].fields[o.fields['name']] = args[2]; return None
                if mn == 'get':
                    return args[1].fields.get(o.fields['name'])
                if mn == 'getName':
                    return o.fields['name']
            if o.cls == 'java/lang/annotation/$Proxy':
                return o.fields[mn]
            if o.cls == 'java/util/zip/CRC32':
                import zlib
                if mn == 'update':
                    data = bytes(args[1].a) if ds == '([B)V' else (bytes([args[1] & 0xFF]) if ds == '(I)V' else bytes(args[1].a[args[2]:args[2] + args[3]]))
                    o.fields['crc'] = zlib.crc32(data, o.fields['crc'])
                    return None
                if mn == 'getValue':
                    return o.fields['crc']
                if mn == 'reset':
                    o.fields['crc'] = 0; return None
            if o.cls == 'java/security/MessageDigest':
                import hashlib
                if mn == 'digest':
                    if ds == '([B)[B':
                        o.fields['h'].update(bytes(args[1].a))
                    d = o.fields['h'].digest()
                    o.fields['h'] = hashlib.new(o.fields['alg'])
                    return JArray('B', bytearray(d))
                if mn == 'update':
                    o.fields['h'].update(bytes(args[1].a) if ds == '([B)V' else bytes(args[1].a[args[2]:args[2] + args[3]]))
                    return None
                if mn == 'reset':
                    o.fields['h'] = hashlib.new(o.fields['alg']); return None
        if isinstance(o, str):
            if mn == 'getClass':
                return ('class', 'java/lang/String')
            if mn == 'length':
                return len(o)
            if mn == 'charAt':
                return ord(o[args[1]])
            if mn == 'equals':
                return 1 if o == args[1] else 0
            if mn == 'hashCode':
                h = 0
                for ch in o:
                    h = i32(31 * h + ord(ch))
                return h
            if mn == 'toString':
                return o
            if mn == 'split':
                import re as _re
                parts = o.split(args[1]) if len(args[1]) == 1 and args[1] not in '.$|()[{^?*+\\' else _re.split(args[1], o)
                while parts and parts[-1] == '':
                    parts.pop()
                return JArray('L', parts)
            if mn == 'toCharArray':
                return JArray('C', [ord(ch) for ch in o])
            if mn == 'getBytes':
                return JArray('B', bytearray(o.encode('latin-1')))
            if mn == 'isEmpty':
                return 1 if not o else 0
            if mn == 'intern':
                return o
            if mn == 'startsWith':
                return 1 if o.startswith(args[1]) else 0
            if mn == 'substring':
                return o[args[1]:] if len(args) == 2 else o[args[1]:args[2]]
        if isinstance(o, JObject):
            if mn == 'ordinal':
                return o.fields['ordinal']
            if mn == 'name' or (mn == 'toString' and 'name' in o.fields and 'ordinal' in o.fields):
                return o.fields['name']
            if 'list' in o.fields and o.cls in JDK_LISTS + ('java/util/List', 'java/util/Collection'):
                lst = o.fields['list']
                if mn == 'add':
                    if ds.startswith('(I'):
                        lst.insert(args[1], args[2]); return None
                    lst.append(args[1]); return 1
                if mn == 'addAll':
                    lst.extend(args[1].fields['list']); return 1
                if mn in ('addLast',):
                    lst.append(args[1]); return None
                if mn == 'get':
                    if args[1] < 0 or args[1] >= len(lst):
                        self.throw('java/lang/IndexOutOfBoundsException', str(args[1]))
                    return lst[args[1]]
                if mn == 'size':
                    return len(lst)
                if mn == 'isEmpty':
                    return 1 if not lst else 0
                if mn == 'clear':
                    del lst[:]; return None
                if mn in ('iterator', 'listIterator'):
                    it = JObject('java/util/Iterator'); it.fields['it'] = lst; it.fields['i'] = 0; return it
                if mn in ('getFirst', 'peekFirst', 'peek'):
                    if not lst and mn == 'getFirst':
                        self.throw('java/util/NoSuchElementException', None)
                    return lst[0] if lst else None
                if mn in ('removeFirst', 'poll', 'pollFirst'):
                    if not lst and mn == 'removeFirst':
                        self.throw('java/util/NoSuchElementException', None)
                    return lst.pop(0) if lst else None
                if mn == 'remove' and ds.startswith('(I'):
                    return lst.pop(args[1])
                if mn == 'contains':
                    k = _key(args[1]); return 1 if any(_key(v) == k for v in lst) else 0
                if mn == 'toArray':
                    return JArray('L', list(lst))
            if o.cls == 'java/util/Iterator':
                if mn == 'hasNext':
                    return 1 if o.fields['i'] < len(o.fields['it']) else 0
                if mn == 'next':
                    v = o.fields['it'][o.fields['i']]; o.fields['i'] += 1; return v
            if 'map' in o.fields and o.cls in JDK_MAPS:
                mp = o.fields['map']
                cmpo = o.fields.get('cmp')

                def find(key):
                    """dictionary key of the entry equal to `key`: by the TreeMap's comparator when it has one"""
                    if cmpo is None:
                        return _key(key)
                    mc = self.find_method(cmpo.cls, 'compare', '(Ljava/lang/Object;Ljava/lang/Object;)I')
                    for dk, (k0, _) in mp.items():
                        if self.invoke(mc, [cmpo, key, k0]) == 0:
                            return dk
                    return _key(key)

                def ordered():
                    items = list(mp.values())
                    if o.cls == 'java/util/TreeMap':
                        import functools
                        if cmpo is not None:
                            mc = self.find_method(cmpo.cls, 'compare', '(Ljava/lang/Object;Ljava/lang/Object;)I')
                            items.sort(key=functools.cmp_to_key(lambda a, b: self.invoke(mc, [cmpo, a[0], b[0]])))
                        else:
                            items.sort(key=lambda kv: (kv[0].fields['value'] if 'value' in kv[0].fields else kv[0].fields['ordinal']) if isinstance(kv[0], JObject) else kv[0])   # boxed numbers / strings, else enums (Enum.compareTo = ordinal)
                    return items
                if mn == 'put':
                    k = find(args[1]); old = mp.get(k, (None, None))
                    mp[k] = (old[0] if old[0] is not None and cmpo is not None else args[1], args[2]); return old[1]
                if mn == 'get':
                    return mp.get(find(args[1]), (None, None))[1]
                if mn == 'containsKey':
                    return 1 if find(args[1]) in mp else 0
                if mn == 'values' and o.cls == 'java/util/TreeMap':
                    r = JObject('java/util/ArrayList'); r.fields['list'] = [v for _, v in ordered()]; return r
                if mn == 'keySet' and o.cls == 'java/util/TreeMap':
                    r = JObject('java/util/ArrayList'); r.fields['list'] = [k for k, _ in ordered()]; return r
                if mn == 'size':
                    return len(mp)
                if mn == 'isEmpty':
                    return 1 if not mp else 0
                if mn == 'clear':
                    mp.clear(); return None
                if mn == 'values':
                    r = JObject('java/util/ArrayList'); r.fields['list'] = [v for _, v in _hash_order(o, mp)]; return r
                if mn == 'keySet':
                    r = JObject('java/util/ArrayList'); r.fields['list'] = [k for k, _ in _hash_order(o, mp)]; return r
            if 'map' in o.fields and o.cls in JDK_SETS:
                mp = o.fields['map']
                if mn == 'add':
                    k = _key(args[1]); had = k in mp; mp[k] = args[1]; return 0 if had else 1
                if mn == 'contains':
                    return 1 if _key(args[1]) in mp else 0
                if mn == 'isEmpty':
                    return 1 if not mp else 0
                if mn == 'size':
                    return len(mp)
                if mn == 'iterator':
                    it = JObject('java/util/Iterator'); it.fields['it'] = list(mp.values()); it.fields['i'] = 0; return it
            if o.cls.startswith('java/lang/') and 'value' in o.fields and mn in (
                    'intValue', 'charValue', 'byteValue', 'shortValue', 'longValue', 'booleanValue'):
                return o.fields['value']
            if o.cls in ('java/lang/StringBuffer',):
                if mn == 'append':
                    o.fields['s'] += self.to_jstring(args[1], ds); return o
                if mn == 'toString':
                    return o.fields['s']
            if o.cls == 'java/lang/StringBuilder':
                if mn == 'append':
                    o.fields['s'] += self.to_jstring(args[1], ds); return o
                if mn == 'toString':
                    return o.fields['s']
            if mn == 'getMessage':
                return o.fields.get('message')
            if mn == 'hashCode':
                return id(o) & 0x7FFFFFFF
            if mn == 'equals':
                return 1 if o is args[1] else 0
            if mn == 'getClass':
                return ('class', o.cls)
            if o.cls == 'java/io/PrintStream':
                return None
            if o.cls == 'java/lang/Integer' and mn == 'intValue':
                return o.fields['value']
        raise NotImplementedError('virtual %s.%s%s on %r' % (cn, mn, ds, type(o)))

    def bb_call(self, b, mn, ds, a):
        arr, off = b.arr, b.offset
        if mn == 'get':
            if ds == '()B':
                b._chk(self, 1); v = arr[off + b.position]; b.position += 1; return i8(v)
            if ds == '(I)B':
                if a[0] < 0 or a[0] >= b.limit:
                    self.throw('java/lang/IndexOutOfBoundsException', None)
                return i8(arr[off + a[0]])
            if ds == '([B)Ljava/nio/ByteBuffer;':
                dst = a[0]; n = len(dst.a)
                b._chk(self, n); dst.a[0:n] = arr[off + b.position: off + b.position + n]; b.position += n; return b
            if ds == '([BII)Ljava/nio/ByteBuffer;':
                dst, dp, n = a
                b._chk(self, n); dst.a[dp:dp + n] = arr[off + b.position: off + b.position + n]; b.position += n; return b
        if mn == 'put':
            if ds == '(B)Ljava/nio/ByteBuffer;':
                b._chkw(self, 1); arr[off + b.position] = a[0] & 0xFF; b.position += 1; return b
            if ds == '(IB)Ljava/nio/ByteBuffer;':
                if a[0] < 0 or a[0] >= b.limit:
                    self.throw('java/lang/IndexOutOfBoundsException', None)
                arr[off + a[0]] = a[1] & 0xFF; return b
            if ds == '([B)Ljava/nio/ByteBuffer;':
                src = a[0]; n = len(src.a)
                b._chkw(self, n); arr[off + b.position: off + b.position + n] = src.a; b.position += n; return b
            if ds == '([BII)Ljava/nio/ByteBuffer;':
                src, sp, n = a
                b._chkw(self, n); arr[off + b.position: off + b.position + n] = src.a[sp:sp + n]; b.position += n; return b
            if ds == '(Ljava/nio/ByteBuffer;)Ljava/nio/ByteBuffer;':
                s = a[0]; n = s.limit - s.position
                b._chkw(self, n)
                arr[off + b.position: off + b.position + n] = s.arr[s.offset + s.position: s.offset + s.limit]
                b.position += n; s.position = s.limit; return b
        if mn == 'getInt':
            fmt = '<i' if b.order is LITTLE else '>i'
            if ds == '()I':
                b._chk(self, 4); v = struct.unpack_from(fmt, arr, off + b.position)[0]; b.position += 4; return v
            if ds == '(I)I':
                if a[0] < 0 or a[0] + 4 > b.limit:
                    self.throw('java/lang/IndexOutOfBoundsException', None)
                return struct.unpack_from(fmt, arr, off + a[0])[0]
        if mn == 'putInt':
            fmt = '<i' if b.order is LITTLE else '>i'
            if ds == '(I)Ljava/nio/ByteBuffer;':
                b._chkw(self, 4); struct.pack_into(fmt, arr, off + b.position, i32(a[0])); b.position += 4; return b
            if ds == '(II)Ljava/nio/ByteBuffer;':
                if a[0] < 0 or a[0] + 4 > b.limit:
                    self.throw('java/lang/IndexOutOfBoundsException', None)
                struct.pack_into(fmt, arr, off + a[0], i32(a[1])); return b
        if mn in ('putShort', 'putLong') and ds in ('(S)Ljava/nio/ByteBuffer;', '(J)Ljava/nio/ByteBuffer;'):
            n, c = (2, 'h') if mn == 'putShort' else (8, 'q')
            v = a[0] & ((1 << (8 * n)) - 1)
            v = v - (1 << (8 * n)) if v >> (8 * n - 1) else v
            b._chkw(self, n); struct.pack_into(('<' if b.order is LITTLE else '>') + c, arr, off + b.position, v); b.position += n; return b
        if mn == 'position':
            if ds == '()I':
                return b.position
            if a[0] < 0 or a[0] > b.limit:
                self.throw('java/lang/IllegalArgumentException', None)
            b.position = a[0]; return b
        if mn == 'limit':
            if ds == '()I':
                return b.limit
            if a[0] < 0 or a[0] > b.cap:
                self.throw('java/lang/IllegalArgumentException', None)
            b.limit = a[0]
            if b.position > b.limit:
                b.position = b.limit
            return b
        if mn == 'remaining':
            return b.limit - b.position
        if mn == 'hasRemaining':
            return 1 if b.position < b.limit else 0
        if mn == 'capacity':
            return b.cap
        if mn == 'order':
            if ds == '()Ljava/nio/ByteOrder;':
                return b.order
            b.order = a[0]; return b
        if mn == 'rewind':
            b.position = 0; return b
        if mn == 'flip':
            b.limit = b.position; b.position = 0; return b
        if mn == 'clear':
            b.position = 0; b.limit = b.cap; return b
        if mn == 'slice':
            n = ByteBuffer(arr, off + b.position, b.limit - b.position)
            n.jarray = b.jarray    # order resets to BIG_ENDIAN, as in Java
            return n
        if mn == 'duplicate':
            n = ByteBuffer(arr, off, b.cap); n.position = b.position; n.limit = b.limit; n.jarray = b.jarray
            return n
        if mn == 'array':
            return b.jarray
        if mn == 'arrayOffset':
            return off
        if mn == 'hasArray':
            return 1
        raise NotImplementedError('ByteBuffer.%s%s' % (mn, ds))


JDK_LISTS = ('java/util/ArrayList', 'java/util/LinkedList')
JDK_MAPS = ('java/util/HashMap', 'java/util/TreeMap', 'java/util/LinkedHashMap')
JDK_SETS = ('java/util/HashSet', 'java/util/TreeSet', 'java/util/LinkedHashSet', 'java/util/EnumSet')


def _hash_order(o, mp):
    """(key, value) pairs in the order a java.util.HashMap iterates them, for maps whose keys are all java.lang.Integer
    (content id -> block / compressor: Slice.external, CompressionHeader.externalCompressors): bucket by bucket,
    bucket = (h ^ h >>> 16) & (capacity - 1) with h = the int itself, capacity = 16 doubled while size > 0.75 capacity
    (a map that only grew), insertion order inside a bucket.  Other maps keep insertion order (nothing pinned reads them
    in order)."""
    items = list(mp.values())
    if o.cls != 'java/util/HashMap' or not items or not all(isinstance(k, JObject) and k.cls == 'java/lang/Integer' for k, _ in items):
        return items
    cap = 16
    while len(items) > 0.75 * cap:
        cap *= 2

    def bucket(k):
        h = k.fields['value'] & 0xFFFFFFFF
        return (h ^ (h >> 16)) & (cap - 1)
    return sorted(items, key=lambda kv: bucket(kv[0]))          # stable: insertion order inside a bucket


def _key(v):
    """dictionary key for a Java object used in a HashMap / HashSet"""
    if isinstance(v, JObject) and 'value' in v.fields and v.cls.startswith('java/lang/'):
        return (v.cls, v.fields['value'])
    if isinstance(v, (str, int, float)) or v is None:
        return v
    return id(v)


class L(int):
    """marker for category-2 long values on the operand stack"""
    __slots__ = ()


class D(float):
    """marker for category-2 double values on the operand stack"""
    __slots__ = ()


WIDE = (L, D)
_DEFAULT = object()

JDK_SUPERS = {
    'java/lang/Object': None,
    'java/lang/Throwable': 'java/lang/Object',
    'java/lang/Exception': 'java/lang/Throwable',
    'java/lang/Error': 'java/lang/Throwable',
    'java/lang/LinkageError': 'java/lang/Error',
    'java/lang/IncompatibleClassChangeError': 'java/lang/LinkageError',
    'java/lang/NoSuchFieldError': 'java/lang/IncompatibleClassChangeError',
    'java/lang/RuntimeException': 'java/lang/Exception',
    'java/lang/IllegalArgumentException': 'java/lang/RuntimeException',
    'java/lang/IllegalStateException': 'java/lang/RuntimeException',
    'java/lang/ArithmeticException': 'java/lang/RuntimeException',
    'java/lang/NullPointerException': 'java/lang/RuntimeException',
    'java/lang/NegativeArraySizeException': 'java/lang/RuntimeException',
    'java/lang/IndexOutOfBoundsException': 'java/lang/RuntimeException',
    'java/lang/ArrayIndexOutOfBoundsException': 'java/lang/IndexOutOfBoundsException',
    'java/nio/BufferUnderflowException': 'java/lang/RuntimeException',
    'java/util/NoSuchElementException': 'java/lang/RuntimeException',
    'java/nio/BufferOverflowException': 'java/lang/RuntimeException',
}

REFERENCE_JAR = '/root/reference/cramtools-3.0.jar'


def disassemble_refs(vm, cname):
    """names of every member the class references -- handy when extending the JDK emulation"""
    cf = vm.load(cname) if not vm.is_jdk(cname) else None
    out = set()
    for i, e in enumerate(cf.cp):
        if e and e[0] in ('field', 'method', 'imethod'):
            out.add((e[0],) + cf.member_ref(i))
    return sorted(out)


# ----------------------------------------------------------------------------- disassembler (for reading the jar)

_OPNAMES = {
    0: 'nop', 1: 'aconst_null', 16: 'bipush', 17: 'sipush', 18: 'ldc', 19: 'ldc_w', 20: 'ldc2_w',
    21: 'iload', 22: 'lload', 23: 'fload', 24: 'dload', 25: 'aload', 46: 'iaload', 47: 'laload', 50: 'aaload',
    51: 'baload', 52: 'caload', 53: 'saload', 54: 'istore', 55: 'lstore', 58: 'astore', 79: 'iastore', 80: 'lastore',
    83: 'aastore', 84: 'bastore', 87: 'pop', 88: 'pop2', 89: 'dup', 90: 'dup_x1', 91: 'dup_x2', 92: 'dup2', 95: 'swap',
    96: 'iadd', 97: 'ladd', 100: 'isub', 101: 'lsub', 104: 'imul', 105: 'lmul', 107: 'dmul', 108: 'idiv', 109: 'ldiv',
    111: 'ddiv', 112: 'irem', 116: 'ineg', 120: 'ishl', 121: 'lshl', 122: 'ishr', 123: 'lshr', 124: 'iushr', 125: 'lushr',
    126: 'iand', 127: 'land', 128: 'ior', 129: 'lor', 130: 'ixor', 132: 'iinc', 133: 'i2l', 135: 'i2d', 136: 'l2i',
    138: 'l2d', 142: 'd2i', 145: 'i2b', 146: 'i2c', 147: 'i2s', 148: 'lcmp', 153: 'ifeq', 154: 'ifne', 155: 'iflt',
    156: 'ifge', 157: 'ifgt', 158: 'ifle', 159: 'if_icmpeq', 160: 'if_icmpne', 161: 'if_icmplt', 162: 'if_icmpge',
    163: 'if_icmpgt', 164: 'if_icmple', 165: 'if_acmpeq', 166: 'if_acmpne', 167: 'goto', 170: 'tableswitch',
    171: 'lookupswitch', 172: 'ireturn', 173: 'lreturn', 176: 'areturn', 177: 'return', 178: 'getstatic',
    179: 'putstatic', 180: 'getfield', 181: 'putfield', 182: 'invokevirtual', 183: 'invokespecial', 184: 'invokestatic',
    185: 'invokeinterface', 187: 'new', 188: 'newarray', 189: 'anewarray', 190: 'arraylength', 191: 'athrow',
    192: 'checkcast', 193: 'instanceof', 196: 'wide', 197: 'multianewarray', 198: 'ifnull', 199: 'ifnonnull',
}
for _i, _n in enumerate(['iconst_m1', 'iconst_0', 'iconst_1', 'iconst_2', 'iconst_3', 'iconst_4', 'iconst_5', 'lconst_0',
                         'lconst_1', 'fconst_0', 'fconst_1', 'fconst_2', 'dconst_0', 'dconst_1']):
    _OPNAMES[2 + _i] = _n
for _b, _p in ((26, 'iload_'), (30, 'lload_'), (34, 'fload_'), (38, 'dload_'), (42, 'aload_'), (59, 'istore_'),
               (63, 'lstore_'), (67, 'fstore_'), (71, 'dstore_'), (75, 'astore_')):
    for _k in range(4):
        _OPNAMES[_b + _k] = _p + str(_k)


def disassemble(cf, name, desc=None):
    """javap -c style listing of one method (with source line numbers) -- used to read htsjdk's bytecode"""
    out = []
    for (nm, ds), m in cf.methods.items():
        if nm != name or (desc and ds != desc) or m.code is None:
            continue
        out.append('%s%s' % (nm, ds))
        code = m.code
        pc = 0
        up = struct.unpack_from
        while pc < len(code):
            op = code[pc]
            nmn = _OPNAMES.get(op, 'op%d' % op)
            arg = ''
            ln = 1
            if op in (16, 188):
                arg = str(i8(code[pc + 1]) if op == 16 else code[pc + 1]); ln = 2
            elif op == 17:
                arg = str(up('>h', code, pc + 1)[0]); ln = 3
            elif op == 18 or op in (19, 20):
                idx = code[pc + 1] if op == 18 else up('>H', code, pc + 1)[0]
                e = cf.cp[idx]
                arg = repr(cf.utf(e[1])) if e[0] in ('string', 'class') else repr(e[1])
                ln = 2 if op == 18 else 3
            elif 21 <= op <= 25 or 54 <= op <= 58:
                arg = str(code[pc + 1]); ln = 2
            elif op == 132:
                arg = '%d %d' % (code[pc + 1], i8(code[pc + 2])); ln = 3
            elif 153 <= op <= 167 or op in (198, 199):
                arg = '-> %d' % (pc + up('>h', code, pc + 1)[0]); ln = 3
            elif op in (178, 179, 180, 181, 182, 183, 184):
                arg = '%s.%s %s' % cf.member_ref(up('>H', code, pc + 1)[0]); ln = 3
            elif op == 185:
                arg = '%s.%s %s' % cf.member_ref(up('>H', code, pc + 1)[0]); ln = 5
            elif op in (187, 189, 192, 193):
                arg = cf.class_name(up('>H', code, pc + 1)[0]); ln = 3
            elif op == 197:
                arg = cf.class_name(up('>H', code, pc + 1)[0]) + ' dims=%d' % code[pc + 3]; ln = 4
            elif op == 170:
                p = (pc + 4) & ~3
                dflt, lo, hi = up('>iii', code, p)
                arg = 'default->%d ' % (pc + dflt) + ' '.join('%d->%d' % (lo + k, pc + up('>i', code, p + 12 + 4 * k)[0]) for k in range(hi - lo + 1))
                ln = p + 12 + 4 * (hi - lo + 1) - pc
            elif op == 171:
                p = (pc + 4) & ~3
                dflt, n = up('>ii', code, p)
                arg = 'default->%d ' % (pc + dflt) + ' '.join('%d->%d' % (up('>i', code, p + 8 + 8 * k)[0], pc + up('>i', code, p + 12 + 8 * k)[0]) for k in range(n))
                ln = p + 8 + 8 * n - pc
            elif op == 196:
                ln = 6 if code[pc + 1] == 132 else 4
                arg = str(up('>H', code, pc + 2)[0])
            out.append('  %4d [L%s] %-16s %s' % (pc, m.line_of(pc), nmn, arg))
            pc += ln
    return '\n'.join(out)
