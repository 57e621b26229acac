// PiperCUDARuntime - source-compatible stand-in for PiperMetalRuntime
// (ocrickard/piper-swift Sources/PiperMetal/PiperMetalRuntime.swift:10-122) on top of the C ABI of
// include/piper_b200.h.  It lives INSIDE the reference's SwiftPM package as one more target (INTEGRATION.md section 3):
// it imports the reference's own PiperCore (PiperAudioFormat / PiperAudioChunk / PiperConfig, PiperAudio.swift:3-25,
// PiperConfig.swift:3-48) and PiperONNX (ONNXModel, ONNXIR.swift:23-93) instead of redeclaring them, so that
// WavFileWriter / AudioPlayer / PiperCLI take its chunks unchanged.  Same stored properties (model, config, options),
// same three initialisers, same synthesize / synthesizeStream signatures and defaults; batch / speaker forms are additive.
// AUTHORED, NOT BUILT in this repository's image (no Swift toolchain) - see INTEGRATION.md.
import CPiperB200
import Foundation
import PiperCore
import PiperONNX

public enum PiperCUDAError: Error, CustomStringConvertible {
    case failed(code: Int32, message: String)
    public var description: String {
        switch self { case let .failed(code, message): return "piper_b200 error \(code): \(message)" }
    }
}

public final class PiperCUDARuntime: @unchecked Sendable {
    public struct Options: Sendable {                     // PiperMetalRuntime.Options, PiperMetalRuntime.swift:11-26
        public var seed: UInt64
        public var preferCPUConv: Bool                    // accepted for source compatibility; there is no CPU path
        public var preferCPUElementwise: Bool
        public var device: Int32 = 0                      // additive
        public var precision: Int32 = PIPER_B200_PRECISION_F16
        public init(seed: UInt64 = 1234) {                // :15-19
            self.seed = seed
            self.preferCPUConv = false
            self.preferCPUElementwise = false
        }
        public init(seed: UInt64 = 1234, preferCPUConv: Bool, preferCPUElementwise: Bool) {   // :21-25
            self.seed = seed
            self.preferCPUConv = preferCPUConv
            self.preferCPUElementwise = preferCPUElementwise
        }
    }

    public let model: ONNXModel                           // :28-30
    public let config: PiperConfig
    public let options: Options
    private var ctx: OpaquePointer?
    private let lock = NSLock()                           // one context = one stream: calls are serialised
    public private(set) var info = piper_b200_voice_info()
    /// Route B cannot throw from the reference's non-throwing init(model:config:options:); a load failure is kept
    /// here and thrown by the first synthesize call (the reference fails at the same place: its GraphExecutor
    /// validates the graph lazily, GraphExecutor.swift:156-200).
    public private(set) var loadError: Error?

    /// init(modelURL:configURL:options:) - PiperMetalRuntime.swift:37-41.  Like the reference it parses with PiperONNX /
    /// PiperCore and delegates to init(model:config:options:) (route B).  `useLibraryLoader: true` takes route A
    /// instead: the library reads both files itself (csrc/onnx_reader.cpp, csrc/piper_config.cpp).
    public convenience init(modelURL: URL, configURL: URL, options: Options = .init()) throws {
        let model = try ONNXModel(modelURL: modelURL)
        let config = try PiperConfig.load(from: configURL)
        self.init(model: model, config: config, options: options)
        if let e = loadError { throw e }
    }

    /// init(model:config:options:) - PiperMetalRuntime.swift:54-59, route B of the C ABI: every initializer goes down
    /// with piper_b200_set_weight, every node with piper_b200_add_node (the binder follows Conv / ConvTranspose /
    /// Gather / Sub / Mul / Exp nodes to find weights a real export leaves anonymous), the config scalars with
    /// piper_b200_set_config; piper_b200_finalize recognises the VITS graph and repacks.
    public init(model: ONNXModel, config: PiperConfig, options: Options = .init()) {
        self.model = model
        self.config = config
        self.options = options
        do {
            var c: OpaquePointer?
            try Self.check(piper_b200_create(options.device, &c), nil)
            ctx = c
            for (name, t) in model.ir.graph.initializers { try Self.push(c, name: name, tensor: t) }
            for node in model.ir.graph.nodes {
                if node.opType == "Constant" {            // its value travels as a weight under the node's output name
                    for a in node.attributes {
                        if case let .tensor(t) = a.kind, let out = node.outputs.first { try Self.push(c, name: out, tensor: t) }
                    }
                    continue
                }
                var stride: Int32 = 1, dilation: Int32 = 1, group: Int32 = 1, padB: Int32 = -1, padE: Int32 = -1
                for a in node.attributes {
                    switch (a.name, a.kind) {
                    case let ("strides", .ints(v)): stride = Int32(v.first ?? 1)
                    case let ("dilations", .ints(v)): dilation = Int32(v.first ?? 1)
                    case let ("group", .int(v)): group = Int32(v)
                    case let ("pads", .ints(v)):
                        padB = Int32(v.first ?? 0)
                        padE = Int32(v.count > 1 ? v[1] : (v.first ?? 0))
                    default: break
                    }
                }
                try Self.withCStrings(node.inputs) { ins in
                    try Self.withCStrings(node.outputs) { outs in
                        try Self.check(piper_b200_add_node(c, node.opType, node.name, ins, Int32(node.inputs.count), outs,
                                                           Int32(node.outputs.count), stride, dilation, group, padB, padE), c)
                    }
                }
            }
            try Self.check(piper_b200_set_config(c, Int32(config.audio.sample_rate), Int32(config.num_speakers),
                                                 config.inference.noise_scale, config.inference.length_scale,
                                                 config.inference.noise_w), c)
            try Self.check(piper_b200_finalize(c), c)
            try Self.check(piper_b200_set_precision(c, options.precision), c)
            try Self.check(piper_b200_get_voice_info(c, &info), c)
        } catch {
            loadError = error
            if let c = ctx { piper_b200_destroy(c); ctx = nil }
        }
    }

    /// Route A (additive): the library's own C++ readers load the two files; `model` / `config` are still parsed with
    /// the reference's types so that the stored properties exist.
    public static func loadWithLibraryReader(modelURL: URL, configURL: URL, options: Options = .init()) throws -> PiperCUDARuntime {
        let rt = PiperCUDARuntime(shell: try ONNXModel(modelURL: modelURL), config: try PiperConfig.load(from: configURL),
                                  options: options)
        var c: OpaquePointer?
        try check(piper_b200_create(options.device, &c), nil)
        rt.ctx = c
        try check(piper_b200_load_voice(c, modelURL.path, configURL.path), c)
        try check(piper_b200_set_precision(c, options.precision), c)
        try check(piper_b200_get_voice_info(c, &rt.info), c)
        return rt
    }

    private init(shell model: ONNXModel, config: PiperConfig, options: Options) {
        self.model = model
        self.config = config
        self.options = options
    }

    deinit { if let c = ctx { piper_b200_destroy(c) } }

    private static func check(_ rc: Int32, _ c: OpaquePointer?) throws {
        if rc != PIPER_B200_OK {
            throw PiperCUDAError.failed(code: rc, message: String(cString: piper_b200_last_error(c)))
        }
    }

    /// ONNXTensor (ONNXIR.swift:23-54) -> piper_b200_set_weight: raw_data when present, else the typed fallback arrays
    private static func push(_ c: OpaquePointer?, name: String, tensor t: ONNXTensor) throws {
        let dtype: Int32
        switch t.dataType {
        case .float: dtype = PIPER_B200_DTYPE_F32
        case .int32: dtype = PIPER_B200_DTYPE_I32
        case .int64: dtype = PIPER_B200_DTYPE_I64
        case .bool: dtype = PIPER_B200_DTYPE_BOOL
        default: return                                    // other dtypes never reach the Piper-VITS binder
        }
        let dims = t.dims.map(Int64.init)
        func send(_ p: UnsafeRawPointer?, _ n: Int) throws {
            try dims.withUnsafeBufferPointer { dp in
                try check(piper_b200_set_weight(c, name, dtype, dp.baseAddress, Int32(dims.count), p, n), c)
            }
        }
        if let raw = t.rawData, !raw.isEmpty {
            try raw.withUnsafeBytes { try send($0.baseAddress, raw.count) }
        } else if !t.floatData.isEmpty {
            try t.floatData.withUnsafeBytes { try send($0.baseAddress, $0.count) }
        } else if !t.int64Data.isEmpty {
            try t.int64Data.withUnsafeBytes { try send($0.baseAddress, $0.count) }
        } else {
            try send(nil, 0)
        }
    }

    private static func withCStrings<R>(_ strings: [String], _ body: (UnsafePointer<UnsafePointer<CChar>?>?) throws -> R) rethrows -> R {
        let dup = strings.map { strdup($0) }
        defer { dup.forEach { free($0) } }
        let ptrs = dup.map { UnsafePointer<CChar>($0) }
        return try ptrs.withUnsafeBufferPointer { try body($0.baseAddress) }
    }

    public var lastRunTimings: piper_b200_timings {       // lastRunTimings, PiperMetalRuntime.swift:33-35
        var t = piper_b200_timings()
        _ = piper_b200_get_timings(ctx, &t)
        return t
    }

    /// synthesize(phonemeIDs:noiseScale:lengthScale:noiseW:) - PiperMetalRuntime.swift:62-80
    public func synthesize(phonemeIDs: [Int], noiseScale: Float? = nil, lengthScale: Float? = nil,
                           noiseW: Float? = nil, speakerID: Int? = nil) throws -> [Float] {
        return try synthesizeBatch(phonemeIDs: [phonemeIDs], noiseScale: noiseScale, lengthScale: lengthScale,
                                   noiseW: noiseW, speakerIDs: speakerID.map { [$0] })[0]
    }

    /// idsFromIPA (PiperCLI.swift:115-136) / ESpeakPhonemizer.phonemeIDs (ESpeakPhonemizer.swift:73-100)
    public func phonemeIDs(ipa: String) throws -> [Int] {
        if let e = loadError { throw e }
        var n: Int64 = 0
        var buf = [Int64](repeating: 0, count: 2 * ipa.utf8.count + 2)
        try Self.check(piper_b200_ipa_to_ids(ctx, ipa, &buf, Int64(buf.count), &n), ctx)
        return buf.prefix(Int(n)).map(Int.init)
    }

    /// additive: a ragged batch of utterances in one call (BASELINE configs 3 and 5)
    public func synthesizeBatch(phonemeIDs: [[Int]], noiseScale: Float? = nil, lengthScale: Float? = nil,
                                noiseW: Float? = nil, speakerIDs: [Int]? = nil) throws -> [[Float]] {
        if let e = loadError { throw e }
        lock.lock()
        defer { lock.unlock() }
        let ids = phonemeIDs.flatMap { $0.map(Int64.init) }
        let lengths = phonemeIDs.map { Int64($0.count) }
        let sids = speakerIDs?.map(Int64.init)
        var res = piper_b200_result()
        try ids.withUnsafeBufferPointer { idp in
            try lengths.withUnsafeBufferPointer { lp in
                var rq = piper_b200_request()
                rq.phoneme_ids = idp.baseAddress
                rq.lengths = lp.baseAddress
                rq.batch = Int32(phonemeIDs.count)
                rq.noise_scale = noiseScale ?? .nan        // NaN -> voice default (PiperMetalRuntime.swift:63-65)
                rq.length_scale = lengthScale ?? .nan
                rq.noise_w = noiseW ?? .nan
                rq.seed = options.seed
                if let s = sids {
                    try s.withUnsafeBufferPointer { sp in
                        rq.speaker_ids = sp.baseAddress
                        try Self.check(piper_b200_synthesize(ctx, &rq, &res), ctx)
                    }
                } else {
                    try Self.check(piper_b200_synthesize(ctx, &rq, &res), ctx)
                }
            }
        }
        var out: [[Float]] = []
        for b in 0..<phonemeIDs.count {
            let lo = Int(res.sample_offsets[b]), hi = Int(res.sample_offsets[b + 1])
            out.append(Array(UnsafeBufferPointer(start: res.audio + lo, count: hi - lo)))
        }
        return out
    }

    /// synthesizeStream(...) - PiperMetalRuntime.swift:84-121, same shape: a detached task feeds the continuation and
    /// onTermination cancels it.  The reference chunks the finished waveform; here the generator runs chunk by chunk:
    /// PIPER_B200_FLAG_STREAM runs the front end and the flow once, then every piper_b200_stream_next decodes the next
    /// frames (plus the receptive-field halo).  Concatenated chunks are bit-identical to synthesize(); the first chunk
    /// is available after front end + flow + one window.  Cancellation is checked between chunks (:103).
    public func synthesizeStream(phonemeIDs: [Int], noiseScale: Float? = nil, lengthScale: Float? = nil,
                                 noiseW: Float? = nil, chunkSize: Int = 2048)
        -> AsyncThrowingStream<PiperAudioChunk, Error> {
        let format = PiperAudioFormat(sampleRate: config.audio.sample_rate, channels: 1)
        precondition(chunkSize > 0)
        return AsyncThrowingStream { continuation in
            let task = Task.detached(priority: .userInitiated) { [self] in
                do {
                    if let e = loadError { throw e }
                    lock.lock()
                    defer { lock.unlock() }
                    let ids = phonemeIDs.map { Int64($0) }
                    let lengths = [Int64(ids.count)]
                    var res = piper_b200_result()
                    try ids.withUnsafeBufferPointer { idp in
                        try lengths.withUnsafeBufferPointer { lp in
                            var rq = piper_b200_request()
                            rq.phoneme_ids = idp.baseAddress
                            rq.lengths = lp.baseAddress
                            rq.batch = 1
                            rq.noise_scale = noiseScale ?? .nan
                            rq.length_scale = lengthScale ?? .nan
                            rq.noise_w = noiseW ?? .nan
                            rq.seed = options.seed
                            rq.flags = UInt32(PIPER_B200_FLAG_STREAM)
                            try Self.check(piper_b200_synthesize(ctx, &rq, &res), ctx)
                        }
                    }
                    let frames = Int64(max(1, (chunkSize + Int(info.hop) - 1) / Int(info.hop)))
                    var pending: [Float] = []
                    var start = 0
                    var isFinal: Int32 = 0
                    while isFinal == 0, !Task.isCancelled {
                        try Self.check(piper_b200_stream_next(ctx, frames, &res, &isFinal), ctx)
                        pending.append(contentsOf: UnsafeBufferPointer(start: res.audio, count: Int(res.total_samples)))
                        while pending.count >= chunkSize && !(isFinal != 0 && pending.count == chunkSize) {
                            continuation.yield(PiperAudioChunk(format: format, startSampleIndex: start,
                                                               samples: Array(pending[0..<chunkSize]), isFinal: false))
                            pending.removeFirst(chunkSize)
                            start += chunkSize
                        }
                    }
                    if !pending.isEmpty, !Task.isCancelled {
                        continuation.yield(PiperAudioChunk(format: format, startSampleIndex: start, samples: pending,
                                                           isFinal: true))
                    }
                    continuation.finish()
                } catch { continuation.finish(throwing: error) }
            }
            continuation.onTermination = { _ in
                task.cancel()
            }
        }
    }
}
