/**
 * stitch-b200.ts -- TypeScript glue between image-stitch's host (layout, z-order, clipping, inflate/deflate,
 * container writing: src/image-concat-core.ts, src/positioned-layout.ts) and libstitch_b200 through the N-API addon
 * native/stitch_b200_napi.cc.  A drop-in for the pixel path only: the generators generateFilteredScanlines
 * (image-concat-core.ts:389-545) / generatePositionedScanlines (:551-685) + filterScanline (:1276-1303) become
 * `filteredBands()`, the strip feed of streamJpegData (:881-914) becomes `jpegBands()`.
 *
 * NOT compiled or run in this repository's containers (no node, no tsc: probed); the Python host
 * image_stitch_b200/api.py is the tested twin of this file.
 */
import { createRequire } from 'node:module';

// Types of the reference this file plugs into (src/types.ts, src/decoders/types.ts, src/positioned-layout.ts).
export interface PngHeader { width: number; height: number; bitDepth: number; colorType: number; interlaceMethod: number }
export interface ImageDecoderLike {
  getHeader(): Promise<{ width: number; height: number; channels: number; bitDepth: number }>;
  scanlines(): AsyncGenerator<Uint8Array>;
  close(): Promise<void>;
  /** Optional: the inflated, still filtered IDAT stream -- the path then starts at kernel (a) on the device. */
  inflated?(): Promise<{ data: Uint8Array; header: PngHeader }>;
  /** Optional: the concatenated IDAT payloads (zlib stream): inflated on the device (filtered = 2). */
  deflated?(): Promise<{ data: Uint8Array; header: PngHeader }>;
  /** Optional (JPEG decoders): the file itself -- decoded on the device by stitch_job_add_tile_jpeg. */
  jpegFile?(): Uint8Array;
}
export interface PlacedTile {
  imageIdx: number; x: number; y: number;           // clipped top-left on the canvas (>= 0)
  sourceOffsetX: number; sourceOffsetY: number;     // ClippedImageInfo (positioned-layout.ts:107-190)
  visibleWidth: number; visibleHeight: number;
  zIndex: number;                                   // given ?? input index (positioned-layout.ts:184)
}
export const enum OutKind { PngFiltered = 0, RawRgba = 1, JpegEntropy = 2, PngDeflate = 3 }

interface Native {
  createContext(device: number): unknown;
  createJob(ctx: unknown, desc: Record<string, unknown>): unknown;
  addTile(job: unknown, desc: Record<string, unknown>, data: Uint8Array, plte?: Uint8Array, trns?: Uint8Array): void;
  addTileJpeg(job: unknown, placement: Record<string, unknown>, file: Uint8Array): void;
  jpegInfo(ctx: unknown, file: Uint8Array): { width: number; height: number; channels: number };
  runBand(job: unknown, y0: number, y1: number, out: Uint8Array): Promise<number>;
  jpegHeader(width: number, height: number, quality: number): Uint8Array;
  destroyJob(job: unknown): void;
  destroyContext(ctx: unknown): void;
}
const native: Native = createRequire(import.meta.url)('../../build/Release/stitch_b200_napi.node');

const colorTypeOfChannels = (channels: number): number => ({ 1: 0, 2: 4, 3: 2, 4: 6 } as Record<number, number>)[channels]; // image-concat-core.ts:47-71

async function drainScanlines(d: ImageDecoderLike): Promise<Uint8Array> {   // foreign decoders: raw rows (filtered = 0)
  const rows: Uint8Array[] = [];
  for await (const r of d.scanlines()) rows.push(r);
  const out = new Uint8Array(rows.reduce((n, r) => n + r.length, 0));
  let o = 0;
  for (const r of rows) { out.set(r, o); o += r.length; }
  return out;
}

export interface GpuJobOptions {
  device?: number;
  canvasWidth: number; canvasHeight: number;
  outBitDepth: 8 | 16;                 // determineCommonFormat (pixel-ops.ts:293-307); 8 for JPEG (:1026-1027)
  outKind: OutKind;
  alphaBlend: boolean;                 // enableAlphaBlending !== false, positioned mode only
  background: Uint8Array;              // getTransparentColor (pixel-ops.ts:255-287): 4 or 8 bytes, output format
  jpegQuality?: number;
  deviceInflate?: boolean;
}

/** One canvas on one device: tiles in, bands of rows out (ascending y; the caller feeds deflate / the file). */
export class GpuStitchJob {
  private ctx: unknown;
  private job: unknown;
  constructor(private readonly o: GpuJobOptions) {
    this.ctx = native.createContext(o.device ?? 0);
    this.job = native.createJob(this.ctx, {
      canvasW: o.canvasWidth, canvasH: o.canvasHeight, outBitDepth: o.outBitDepth, outKind: o.outKind,
      alphaBlend: o.alphaBlend, jpegQuality: o.jpegQuality ?? 85, background: o.background,
    });
  }

  /** Registers one input; errors carry the reference's messages (e.g. "Unsupported source color type: 3"). */
  async addTile(decoder: ImageDecoderLike, t: PlacedTile): Promise<void> {
    let data: Uint8Array; let header: PngHeader; let filtered = 0;
    if (decoder.jpegFile) {   // a JPEG input: the file itself, decoded on the device (SURVEY 8f.4)
      try {
        native.addTileJpeg(this.job, { dstX: t.x, dstY: t.y, srcOffX: t.sourceOffsetX, srcOffY: t.sourceOffsetY,
          visW: t.visibleWidth, visH: t.visibleHeight, zIndex: t.zIndex, imageIdx: t.imageIdx }, decoder.jpegFile());
        return;
      } catch (e) {   // progressive, 12-bit, CMYK ...: the reference's own decoders take it (raw rows below)
        if (!String((e as Error).message).startsWith('Unsupported JPEG')) throw e;
      }
    }
    if (this.o.deviceInflate && decoder.deflated) { ({ data, header } = await decoder.deflated()); filtered = 2; }
    else if (decoder.inflated) { ({ data, header } = await decoder.inflated()); filtered = 1; }
    else {
      const h = await decoder.getHeader();
      header = { width: h.width, height: h.height, bitDepth: h.bitDepth, colorType: colorTypeOfChannels(h.channels), interlaceMethod: 0 };
      data = await drainScanlines(decoder);
    }
    native.addTile(this.job, {
      width: header.width, height: header.height, bitDepth: header.bitDepth, colorType: header.colorType,
      interlace: filtered ? header.interlaceMethod : 0, filtered,
      dstX: t.x, dstY: t.y, srcOffX: t.sourceOffsetX, srcOffY: t.sourceOffsetY, visW: t.visibleWidth, visH: t.visibleHeight,
      zIndex: t.zIndex, imageIdx: t.imageIdx,
    }, data);
  }

  /** Rows of [filter byte + W*bpp] (PNG) or W*bpp (raw), `bandRows` at a time, in file order. */
  async *filteredBands(bandRows: number): AsyncGenerator<Uint8Array> {
    const bpp = this.o.outBitDepth === 16 ? 8 : 4;
    const rowBytes = this.o.canvasWidth * bpp + (this.o.outKind === OutKind.PngFiltered ? 1 : 0);
    const buf = new Uint8Array(bandRows * rowBytes);
    for (let y0 = 0; y0 < this.o.canvasHeight; y0 += bandRows) {
      const y1 = Math.min(this.o.canvasHeight, y0 + bandRows);
      const n = await native.runBand(this.job, y0, y1, buf);
      yield buf.subarray(0, n);
    }
  }

  /** JPEG: header (SOI .. SOS, restart interval = one MCU row), entropy-coded bands (multiples of 8 rows), EOI. */
  async *jpegBands(bandRows: number): AsyncGenerator<Uint8Array> {
    const rows = Math.max(8, bandRows - (bandRows % 8));
    yield native.jpegHeader(this.o.canvasWidth, this.o.canvasHeight, this.o.jpegQuality ?? 85);
    const mcus = Math.ceil(this.o.canvasWidth / 8) * (rows / 8);
    const buf = new Uint8Array(24 + mcus * 3 * 256 + 64);   // stitch_jpeg_band_info + worst case of baseline Huffman coding
    for (let y0 = 0; y0 < this.o.canvasHeight; y0 += rows) {
      const n = await native.runBand(this.job, y0, Math.min(this.o.canvasHeight, y0 + rows), buf);
      yield buf.subarray(24, n).slice();   // the payload after the band header: byte-aligned restart segments
    }
    yield Uint8Array.of(0xff, 0xd9);
  }

  close(): void {
    native.destroyJob(this.job);
    native.destroyContext(this.ctx);
  }
}

/**
 * Where the reference calls this (grid mode, PNG): CoreStreamingConcatenator.streamCompressedData
 * (image-concat-core.ts:309-384) keeps its deflate batching and IDAT writing and replaces
 *   for await (const scanline of this.generateFilteredScanlines(...)) { deflator.push(scanline) ... }
 * by
 *   const job = new GpuStitchJob({...});  for (cell of cells) await job.addTile(decoders[cell.imageIdx], cell);
 *   for await (const band of job.filteredBands(bandRows))
 *     for (let o = 0; o < band.length; o += scanlineSize) { await deflator.push(band.subarray(o, o + scanlineSize)); ...flush every N }
 * onProgress fires when an input's last row lies in a finished band (:486-492); clip warnings stay on the host.
 */
