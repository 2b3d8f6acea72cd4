// UNCOMPILED SOURCE — there is no JDK, Maven or Flink jar in the build image (SURVEY.md §8c), so this
// file has never been through javac.  It is the operator a maintainer of the reference would add to
// replace
//     .keyBy(1).timeWindow(Time.minutes(5), Time.seconds(5)).reduce(..).map(..).filter(..)
// (chapter3/src/main/java/me/zjy/BandwidthMonitorWithEventTime.java:45-55) by
//     .keyBy(1).transform("gpuWindowAgg", outType, new GpuWindowAggOperator(cfg))
// and it follows, call for call, the C++ and Python mirrors that ARE compiled and tested here
// (monitor_systam_flink_quickstart_b200/host/gpu_window_operator.hpp, .../operator.py).
package me.zjy;

import java.nio.ByteBuffer;
import java.nio.ByteOrder;
import java.util.ArrayList;
import java.util.HashMap;
import java.util.List;
import java.util.Map;

import org.apache.flink.api.java.tuple.Tuple3;
import org.apache.flink.streaming.api.operators.AbstractStreamOperator;
import org.apache.flink.streaming.api.operators.OneInputStreamOperator;
import org.apache.flink.streaming.api.watermark.Watermark;
import org.apache.flink.streaming.runtime.streamrecord.StreamRecord;

/** Input (time, channel, flow) as produced by the map at BandwidthMonitorWithEventTime.java:36-45;
 *  output (channel, windowEnd, Mbps) for the windows that pass the alert filter (:48-55). */
public class GpuWindowAggOperator extends AbstractStreamOperator<Tuple3<String, Long, Double>>
        implements OneInputStreamOperator<Tuple3<Integer, String, Long>, Tuple3<String, Long, Double>> {

    private static final long serialVersionUID = 1L;
    static { System.loadLibrary("fmq_b200_jni"); }          // integration/jni/fmq_b200_jni.c over libfmq_b200.so

    // ---- thin natives: one per C ABI entry point used (include/fmq_b200.h) ----
    private static native long nCreate(ByteBuffer cfg);                                                     // fmq_create
    private static native int nPushBatch(long h, ByteBuffer ts, ByteBuffer key, ByteBuffer val, int n, long watermarkAfter);   // fmq_push_batch(.., FMQ_MEM_HOST)
    private static native int nFlush(long h, long watermark);                                               // fmq_flush
    private static native int nPollWindowAlerts(long h, ByteBuffer key, ByteBuffer windowEnd, ByteBuffer value, int cap, long[] nOutLeft);  // fmq_poll_window_alerts
    private static native long nLateDropped(long h);                                                        // fmq_get_counters().late_dropped
    private static native String nLastError(long h);                                                        // fmq_last_error
    private static native void nDestroy(long h);                                                            // fmq_destroy

    private static final long FMQ_WM_KEEP = Long.MIN_VALUE + 1;   // include/fmq_b200.h: no watermark arrives with this batch
    private static final int CAP = 1 << 20;                 // records per micro-batch
    private final byte[] configBytes;                       // a serialized fmq_config (see GpuWindowAggConfig)

    private transient long handle;
    private transient ByteBuffer ts, key, val, outKey, outEnd, outVal;
    private transient int n;
    private transient Map<String, Long> dict;                // String key -> dense id (FMQ_KEYS_DENSE)
    private transient List<String> names;

    public GpuWindowAggOperator(byte[] configBytes) { this.configBytes = configBytes; }

    private static ByteBuffer direct(int bytes) { return ByteBuffer.allocateDirect(bytes).order(ByteOrder.nativeOrder()); }

    @Override
    public void open() throws Exception {
        super.open();
        ByteBuffer cfg = direct(configBytes.length);
        cfg.put(configBytes).flip();
        handle = nCreate(cfg);                               // allocates every device buffer once
        if (handle == 0) throw new IllegalStateException("fmq_create failed: " + nLastError(0));
        ts = direct(8 * CAP); key = direct(8 * CAP); val = direct(8 * CAP);
        outKey = direct(8 * CAP); outEnd = direct(8 * CAP); outVal = direct(8 * CAP);
        dict = new HashMap<>(); names = new ArrayList<>();
        n = 0;
    }

    @Override
    public void processElement(StreamRecord<Tuple3<Integer, String, Long>> record) throws Exception {
        if (n == CAP) pushAndDrain(FMQ_WM_KEEP);             // buffer full between two upstream watermarks: event time does not move
        Tuple3<Integer, String, Long> v = record.getValue();
        Long id = dict.get(v.f1);
        if (id == null) { id = (long) names.size(); dict.put(v.f1, id); names.add(v.f1); }
        ts.putLong(8 * n, record.getTimestamp());
        key.putLong(8 * n, id);
        val.putLong(8 * n, v.f2);
        n++;
    }

    @Override
    public void processWatermark(Watermark mark) throws Exception {
        pushAndDrain(mark.getTimestamp());                   // every buffered record precedes this watermark
        super.processWatermark(mark);                        // forward it downstream AFTER the fired windows
    }

    private void pushAndDrain(long watermarkAfter) {
        check(nPushBatch(handle, ts, key, val, n, watermarkAfter));
        n = 0;
        long[] outLeft = new long[2];
        do {
            check(nPollWindowAlerts(handle, outKey, outEnd, outVal, CAP, outLeft));
            for (int i = 0; i < (int) outLeft[0]; i++) {
                long end = outEnd.getLong(8 * i);
                output.collect(new StreamRecord<>(
                        new Tuple3<>(names.get((int) outKey.getLong(8 * i)), end, outVal.getDouble(8 * i)),
                        end - 1));                           // Flink stamps window results with window.maxTimestamp()
            }
        } while (outLeft[1] > 0);
    }

    private void check(int status) {
        if (status != 0) throw new IllegalStateException("fmq status " + status + ": " + nLastError(handle));
    }

    /** WindowOperator's numLateRecordsDropped. */
    public long numLateRecordsDropped() { return nLateDropped(handle); }

    @Override
    public void close() throws Exception {
        if (handle != 0) {
            if (n > 0) pushAndDrain(FMQ_WM_KEEP);
            check(nFlush(handle, Long.MAX_VALUE));           // end of a finite source: Watermark(Long.MAX_VALUE)
            pushAndDrain(Long.MAX_VALUE);
            nDestroy(handle);
            handle = 0;
        }
        super.close();
    }
}
