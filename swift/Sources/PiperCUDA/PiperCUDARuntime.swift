// PiperCUDARuntime - drop-in for PiperMetalRuntime
// (ocrickard/piper-swift Sources/PiperMetal/PiperMetalRuntime.swift:10-122) on top of the C ABI of
// include/piper_b200.h.  Same initialisers, same synthesize(phonemeIDs:noiseScale:lengthScale:noiseW:)
// signature and defaults, same chunked synthesizeStream; additive batch / speaker forms.
// AUTHORED, NOT BUILT in this repository's image (no Swift toolchain) - see INTEGRATION.md.
import CPiperB200
import Foundation

public enum PiperCUDAError: Error, CustomStringConvertible {
    case failed(code: Int32, message: String)
    public var description: String {
        switch self { case let .failed(code, message): return "piper_b200 error \(code): \(message)" }
    }
}

public struct PiperAudioFormat: Sendable { public let sampleRate: Int; public let channels: Int }   // PiperAudio.swift:3-11
public struct PiperAudioChunk: Sendable {                                                           // PiperAudio.swift:13-25
    public let format: PiperAudioFormat
    public let startSampleIndex: Int
    public let samples: [Float]
    public let isFinal: Bool
}

public final class PiperCUDARuntime: @unchecked Sendable {
    public struct Options: Sendable {                     // PiperMetalRuntime.Options, PiperMetalRuntime.swift:11-26
        public var seed: UInt64 = 1234
        public var preferCPUConv = false                  // accepted for source compatibility; there is no CPU path
        public var preferCPUElementwise = false
        public var device: Int32 = 0
        public var precision: Int32 = PIPER_B200_PRECISION_F16
        public init() {}
    }

    private var ctx: OpaquePointer?
    private let options: Options
    public private(set) var info = piper_b200_voice_info()

    /// init(modelURL:configURL:options:) - PiperMetalRuntime.swift:37-41 (route A: the library parses the files)
    public init(modelURL: URL, configURL: URL, options: Options = Options()) throws {
        self.options = options
        var c: OpaquePointer?
        try Self.check(piper_b200_create(options.device, &c), nil)
        ctx = c
        try Self.check(piper_b200_load_voice(c, modelURL.path, configURL.path), c)
        try Self.check(piper_b200_set_precision(c, options.precision), c)
        try Self.check(piper_b200_get_voice_info(c, &info), c)
    }

    deinit { if let c = ctx { piper_b200_destroy(c) } }

    private static func check(_ rc: Int32, _ c: OpaquePointer?) throws {
        if rc != PIPER_B200_OK {
            throw PiperCUDAError.failed(code: rc, message: String(cString: piper_b200_last_error(c)))
        }
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
        var n: Int64 = 0
        var buf = [Int64](repeating: 0, count: 2 * ipa.utf8.count + 2)
        try Self.check(piper_b200_ipa_to_ids(ctx, ipa, &buf, Int64(buf.count), &n), ctx)
        return buf.prefix(Int(n)).map(Int.init)
    }

    /// additive: a ragged batch of utterances in one call (BASELINE configs 3 and 5)
    public func synthesizeBatch(phonemeIDs: [[Int]], noiseScale: Float? = nil, lengthScale: Float? = nil,
                                noiseW: Float? = nil, speakerIDs: [Int]? = nil) throws -> [[Float]] {
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

    /// synthesizeStream(...) - PiperMetalRuntime.swift:84-121.  The reference chunks the finished waveform; here
    /// the generator runs chunk by chunk: PIPER_B200_FLAG_STREAM runs the front end and the flow once, then every
    /// piper_b200_stream_next decodes the next frames (plus the receptive-field halo).  Concatenated chunks are
    /// bit-identical to synthesize(); the first chunk is available after front end + flow + one window.
    public func synthesizeStream(phonemeIDs: [Int], noiseScale: Float? = nil, lengthScale: Float? = nil,
                                 noiseW: Float? = nil, chunkSize: Int = 2048)
        -> AsyncThrowingStream<PiperAudioChunk, Error> {
        AsyncThrowingStream { continuation in
            do {
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
                        rq.seed = self.options.seed
                        rq.flags = UInt32(PIPER_B200_FLAG_STREAM)
                        try Self.check(piper_b200_synthesize(self.ctx, &rq, &res), self.ctx)
                    }
                }
                let fmt = PiperAudioFormat(sampleRate: Int(info.sample_rate), channels: 1)
                let frames = Int64(max(1, (chunkSize + Int(info.hop) - 1) / Int(info.hop)))
                var pending: [Float] = []
                var start = 0
                var isFinal: Int32 = 0
                while isFinal == 0 {
                    try Self.check(piper_b200_stream_next(self.ctx, frames, &res, &isFinal), self.ctx)
                    pending.append(contentsOf: UnsafeBufferPointer(start: res.audio, count: Int(res.total_samples)))
                    while pending.count >= chunkSize && !(isFinal != 0 && pending.count == chunkSize) {
                        continuation.yield(PiperAudioChunk(format: fmt, startSampleIndex: start,
                                                           samples: Array(pending[0..<chunkSize]), isFinal: false))
                        pending.removeFirst(chunkSize)
                        start += chunkSize
                    }
                }
                if !pending.isEmpty {
                    continuation.yield(PiperAudioChunk(format: fmt, startSampleIndex: start, samples: pending,
                                                       isFinal: true))
                }
                continuation.finish()
            } catch { continuation.finish(throwing: error) }
        }
    }
}
