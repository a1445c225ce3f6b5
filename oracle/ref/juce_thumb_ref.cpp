// oracle/ref/juce_thumb_ref.cpp — TEST INFRASTRUCTURE.  A C wrapper that RUNS the reference's own juce::AudioThumbnail
// (juce_audio_utils/gui/juce_AudioThumbnail.cpp, compiled in place from /root/reference by oracle/Makefile) the way
// WaveFormItem does (lib/WaveFormItem.cpp:22,101-143): AudioFormatManager with the WAV format, setSource(InputSource),
// level 0 filled by the thumbnail's own TimeSliceThread through LevelDataSource::readNextBlock -> readMaxLevels, then
// drawChannel(g, bounds, start, end, 0, 1.0f).
//
// juce_audio_utils as a module does not build here (it pulls juce_gui_basics -> X11/freetype), but AudioThumbnail itself
// is CPU arithmetic on top of juce_core / juce_audio_formats / juce_events and the header-only geometry of juce_graphics.
// It needs three juce_graphics symbols at link time: Colour(uint32) (static colour tables in the headers) and the two
// Graphics members CachedWindow::drawChannel calls.  They are defined below as what juce::Graphics does on the reference's
// own context: forward to the LowLevelGraphicsContext (juce_GraphicsContext.cpp:124-127,366-369), which in libzl is
// QPainterContext (lib/QPainterContext.cpp:96-108 clip = device rectangle, :190-200 fillRectList ->
// QPainter::fillRect(int, int, int, int): the float rectangle is truncated to ints).  Qt itself: unavailable on this box;
// the integer column fill is restated here (four lines) and is the only restated step on this path.
//
// Nothing under libzl_b200/ links, loads or calls this file.
#include <juce_core/juce_core.h>
#include <juce_audio_basics/juce_audio_basics.h>
#include <juce_audio_formats/juce_audio_formats.h>
#include <juce_events/juce_events.h>
#include <juce_graphics/juce_graphics.h>

// the level-0 table (AudioThumbnail::channels -> ThumbData::data) and the pixel cache (CachedWindow::data) are private
#define private public
#include <juce_audio_utils/gui/juce_AudioThumbnailBase.h>
#include <juce_audio_utils/gui/juce_AudioThumbnail.h>
#include <juce_audio_utils/gui/juce_AudioThumbnailCache.h>
#include <juce_audio_utils/gui/juce_AudioThumbnail.cpp>
#include <juce_audio_utils/gui/juce_AudioThumbnailCache.cpp>
#undef private

#include <cstring>
#include <vector>

namespace {
struct Sink {
    juce::Rectangle<int> clip;
    std::vector<juce::Rectangle<float>> rects;
};
Sink* gSink = nullptr;      // one draw at a time (the tests are single-threaded)
}  // namespace

namespace juce {
Colour::Colour (uint32 col) noexcept : argb (static_cast<uint8> ((col >> 24) & 0xff), static_cast<uint8> ((col >> 16) & 0xff),
                                             static_cast<uint8> ((col >> 8) & 0xff), static_cast<uint8> (col & 0xff)) {}
Rectangle<int> Graphics::getClipBounds() const { return gSink ? gSink->clip : Rectangle<int>(); }
void Graphics::fillRectList (const RectangleList<float>& list) const
{
    if (gSink == nullptr) return;
    for (int i = 0; i < list.getNumRectangles(); ++i) gSink->rects.push_back (list.getRectangle (i));
}
}  // namespace juce

namespace {

// the file as WaveFormItem::setSource sees it (FileInputSource, lib/WaveFormItem.cpp:107), held in memory
class MemorySource : public juce::InputSource {
public:
    MemorySource (const void* d, size_t n) : data (d), size (n) {}
    juce::InputStream* createInputStream() override { return new juce::MemoryInputStream (data, size, false); }
    juce::InputStream* createInputStreamFor (const juce::String&) override { return nullptr; }
    juce::int64 hashCode() const override { return (juce::int64) size * 31 + 7; }
private:
    const void* data; size_t size;
};

struct Thumb {
    juce::AudioFormatManager formats;
    juce::AudioThumbnailCache cache { 1 };
    std::unique_ptr<juce::AudioThumbnail> thumb;
    std::vector<unsigned char> image;
};

}  // namespace

extern "C" {

// spt: sourceSamplesPerThumbnailSample (WaveFormItem: 512).  wav: a RIFF/WAVE image; copied.  Blocks until the thumbnail's
// background thread has read the whole file (isFullyLoaded).  Returns a handle or null.
__attribute__((visibility("default"))) void* ref_thumb_create(const unsigned char* wav, long long nbytes, int spt) {
    auto* t = new Thumb;
    t->formats.registerFormat(new juce::WavAudioFormat(), true);
    t->image.assign(wav, wav + nbytes);
    t->thumb.reset(new juce::AudioThumbnail(spt, t->formats, t->cache));
    if (!t->thumb->setSource(new MemorySource(t->image.data(), t->image.size()))) { delete t; return nullptr; }
    for (int i = 0; i < 600000 && !t->thumb->isFullyLoaded(); ++i) juce::Thread::sleep(1);
    if (!t->thumb->isFullyLoaded()) { delete t; return nullptr; }
    return t;
}
__attribute__((visibility("default"))) void ref_thumb_destroy(void* h) { delete static_cast<Thumb*>(h); }

__attribute__((visibility("default"))) int ref_thumb_info(void* h, int* nThumb, int* nch, double* totalLength) {
    auto* t = static_cast<Thumb*>(h);
    if (nThumb) *nThumb = t->thumb->numChannels > 0 ? t->thumb->channels.getUnchecked(0)->getSize() : 0;
    if (nch) *nch = t->thumb->getNumChannels();
    if (totalLength) *totalLength = t->thumb->getTotalLength();
    return 0;
}

// out[nch][nThumb][2]
__attribute__((visibility("default"))) int ref_thumb_level0(void* h, signed char* out) {
    auto* t = static_cast<Thumb*>(h);
    const int nch = t->thumb->getNumChannels();
    for (int ch = 0; ch < nch; ++ch) {
        auto* td = t->thumb->channels.getUnchecked(ch);
        const int n = td->getSize();
        for (int i = 0; i < n; ++i) {
            const auto* mm = td->getData(i);
            out[((size_t) ch * (size_t) n + (size_t) i) * 2] = mm->getMinValue();
            out[((size_t) ch * (size_t) n + (size_t) i) * 2 + 1] = mm->getMaxValue();
        }
    }
    return 0;
}

// CachedWindow::refillCache as drawChannel calls it; out[nch][width][2].  Returns 1, or 0 where refillCache returns false.
__attribute__((visibility("default"))) int ref_thumb_pixels(void* h, double t0, double t1, int width, signed char* out) {
    auto* t = static_cast<Thumb*>(h);
    auto& th = *t->thumb;
    const juce::ScopedLock sl(th.lock);
    th.window->invalidate();
    if (!th.window->refillCache(width, t0, t1, th.sampleRate, th.numChannels, th.samplesPerThumbSample, th.source.get(), th.channels)) return 0;
    for (int ch = 0; ch < th.window->numChannelsCached; ++ch)
        for (int x = 0; x < width; ++x) {
            const auto* mm = th.window->getData(ch, x);
            out[((size_t) ch * (size_t) width + (size_t) x) * 2] = mm->getMinValue();
            out[((size_t) ch * (size_t) width + (size_t) x) * 2 + 1] = mm->getMaxValue();
        }
    return 1;
}

// WaveFormItem::paint (lib/WaveFormItem.cpp:130-139): drawChannel over (0, 0, width, height); mask[height][width].
__attribute__((visibility("default"))) int ref_thumb_draw(void* h, int channel, double t0, double t1, int width, int height, float zoom,
                                                          unsigned char* mask) {
    auto* t = static_cast<Thumb*>(h);
    Sink sink;
    sink.clip = juce::Rectangle<int>(0, 0, width, height);           // QPainterContext::getClipBounds with no clip set: the device rectangle
    gSink = &sink;
    alignas(juce::Graphics) static unsigned char gbuf[sizeof(juce::Graphics)];   // never constructed: the two members above touch no state
    juce::Graphics& g = *reinterpret_cast<juce::Graphics*>(gbuf);
    t->thumb->window->invalidate();
    t->thumb->drawChannel(g, juce::Rectangle<int>(0, 0, width, height), t0, t1, channel, zoom);
    gSink = nullptr;
    memset(mask, 0, (size_t) width * (size_t) height);
    for (const auto& r : sink.rects) {
        // QPainterContext::fillRectList -> QPainter::fillRect(int x, int y, int w, int h, brush)   lib/QPainterContext.cpp:190-200
        const int rx = (int) r.getX(), ry = (int) r.getY(), rw = (int) r.getWidth(), rh = (int) r.getHeight();
        if (rw <= 0 || rh <= 0) continue;
        for (int y = std::max(ry, 0); y < std::min(ry + rh, height); ++y)
            for (int x = std::max(rx, 0); x < std::min(rx + rw, width); ++x) mask[(size_t) y * (size_t) width + (size_t) x] = 1;
    }
    return (int) sink.rects.size();
}

// AudioThumbnail::saveTo  juce_AudioThumbnail.cpp:616-639.  Returns the stream size (copied when it fits).
__attribute__((visibility("default"))) long long ref_thumb_save(void* h, unsigned char* out, long long cap) {
    auto* t = static_cast<Thumb*>(h);
    juce::MemoryOutputStream mo;
    t->thumb->saveTo(mo);
    const long long n = (long long) mo.getDataSize();
    if (out && cap >= n) memcpy(out, mo.getData(), (size_t) n);
    return n;
}

// AudioThumbnail::loadFrom (:593-614) into a fresh thumbnail, then its level-0 table: out[nch][nThumb][2].
// Returns nThumb (0 when loadFrom refuses the stream); nch through *nchOut.
__attribute__((visibility("default"))) int ref_thumb_load(const unsigned char* jatm, long long nbytes, int spt, signed char* out, long long cap, int* nchOut) {
    juce::AudioFormatManager fm;
    juce::AudioThumbnailCache cache(1);
    juce::AudioThumbnail th(spt, fm, cache);
    juce::MemoryInputStream in(jatm, (size_t) nbytes, false);
    if (!th.loadFrom(in)) return 0;
    const int nch = th.getNumChannels();
    if (nchOut) *nchOut = nch;
    const int n = nch > 0 ? th.channels.getUnchecked(0)->getSize() : 0;
    if ((long long) nch * n * 2 > cap) return -n;
    for (int ch = 0; ch < nch; ++ch)
        for (int i = 0; i < n; ++i) {
            const auto* mm = th.channels.getUnchecked(ch)->getData(i);
            out[((size_t) ch * (size_t) n + (size_t) i) * 2] = mm->getMinValue();
            out[((size_t) ch * (size_t) n + (size_t) i) * 2 + 1] = mm->getMaxValue();
        }
    return n;
}

}  // extern "C"
