package org.uu.nl.embedding.nativepath;

import grph.properties.Property;
import org.uu.nl.embedding.compare.CompareGroup;
import org.uu.nl.embedding.util.InMemoryRdfGraph;
import org.uu.nl.embedding.util.config.Configuration;

import java.time.LocalDate;
import java.time.format.DateTimeFormatter;
import java.time.format.DateTimeParseException;
import java.util.*;

/**
 * Drop-in for the compare loop of Rdf2GrphConverter.convert (:127-186): instead of one CompareJob
 * per source literal on a work-stealing pool, the group's literals are described by the numbers the
 * metric reads and compared on the device (ge_similarity_pairs).  The caller keeps the edge-adding
 * half of the loop (:160-171) unchanged and feeds it from {@link Result}.
 *
 * Source only: this image has no JDK, so the class is not compiled here (INTEGRATION.md §6).
 */
public final class NativeSimilarity {

    public static final class Result { public int[] vert, other; public float[] similarity; }

    /** profile(label) as the reference's PreComputed metrics build it (ShingleBased / TokenBased getProfile) */
    public interface Profiler { Map<String, Integer> profile(String label); }

    public static Result compare(long ctx, CompareGroup group, Configuration.SimilarityGroup cfg,
                                 InMemoryRdfGraph g, Profiler profiler) {
        final Property labels = g.getVertexLabelProperty();
        final int[] source = group.source.stream().mapToInt(i -> i).toArray();   // same order as :139-140
        final int[] target = group.target.stream().mapToInt(i -> i).toArray();
        final TreeSet<Integer> all = new TreeSet<>(group.source); all.addAll(group.target);
        final int n = all.size();
        final int[] vert = new int[n]; final Map<Integer, Integer> item = new HashMap<>();
        for (int v : all) { vert[item.size()] = v; item.put(v, item.size()); }
        final String[] label = new String[n];
        final int[] stringId = new int[n]; final byte[] flags = new byte[n];
        final Map<String, Integer> ids = new HashMap<>();
        final int method = cfg.getMethodEnum().ordinal();           // NGRAM_COSINE .. TOKEN_JACCARD = 0..3
        final int geMethod = method <= 3 ? method : method <= 5 ? method + 4 : method - 2;   // JAROWINKLER, LEVENSHTEIN -> 8, 9
        for (int k = 0; k < n; k++) {
            label[k] = labels.getValueAsString(vert[k]);
            stringId[k] = ids.computeIfAbsent(label[k], s -> ids.size());
            if (label[k].isEmpty()) flags[k] |= 1;
            if (label[k].length() >= cfg.getNgram()) flags[k] |= 2;
        }
        long[] profOff = null, epochDay = null, packed = null, charOff = null; int[] profKey = null, profCnt = null, cutSlot = null, numVal = null;
        byte[] numStatus = null; int nCut = 0; char[] chars = null;
        if (geMethod >= 8) {                                        // character metrics: the labels themselves
            final StringBuilder sb = new StringBuilder(); charOff = new long[n + 1];
            for (int k = 0; k < n; k++) { sb.append(label[k]); charOff[k + 1] = sb.length(); }
            chars = sb.toString().toCharArray();
        } else if (geMethod <= 3) {                                        // shingle / token ids, sorted per literal
            final Map<String, Integer> vocab = new HashMap<>();
            final List<int[]> rows = new ArrayList<>(); profOff = new long[n + 1];
            for (int k = 0; k < n; k++) {
                final TreeMap<Integer, Integer> row = new TreeMap<>();
                for (Map.Entry<String, Integer> e : profiler.profile(label[k]).entrySet())
                    row.put(vocab.computeIfAbsent(e.getKey(), s -> vocab.size()), e.getValue());
                for (Map.Entry<Integer, Integer> e : row.entrySet()) rows.add(new int[]{e.getKey(), e.getValue()});
                profOff[k + 1] = rows.size();
            }
            profKey = new int[rows.size()]; profCnt = new int[rows.size()];
            for (int p = 0; p < rows.size(); p++) { profKey[p] = rows.get(p)[0]; profCnt[p] = rows.get(p)[1]; }
        } else if (geMethod == 4) {                                 // NUMERIC: parse at every cut position of the group
            final TreeSet<Integer> cuts = new TreeSet<>(); final int[] hat = new int[n];
            for (int k = 0; k < n; k++) cuts.add(hat[k] = label[k].indexOf('^'));
            final List<Integer> cutList = new ArrayList<>(cuts); nCut = cutList.size();
            cutSlot = new int[n]; numVal = new int[n * nCut]; numStatus = new byte[n * nCut];
            for (int k = 0; k < n; k++) {
                cutSlot[k] = cutList.indexOf(hat[k]);
                for (int c = 0; c < nCut; c++) {
                    final int h = cutList.get(c);
                    if (h > label[k].length()) { numStatus[k * nCut + c] = 2; continue; }
                    try { numVal[k * nCut + c] = Integer.parseInt(h < 0 ? label[k] : label[k].substring(0, h)); }
                    catch (NumberFormatException e) { numStatus[k * nCut + c] = 1; }
                }
            }
        } else {                                                    // DATE_*: LocalDate fields
            final DateTimeFormatter f = cfg.getPattern().equals("iso") ? DateTimeFormatter.BASIC_ISO_DATE : DateTimeFormatter.ofPattern(cfg.getPattern());
            epochDay = new long[n]; packed = new long[n];
            for (int k = 0; k < n; k++) {
                final int h = label[k].indexOf('^');
                try {
                    final LocalDate d = LocalDate.parse(h < 0 ? label[k] : label[k].substring(0, h), f);
                    epochDay[k] = d.toEpochDay(); packed[k] = (d.getYear() * 12L + d.getMonthValue() - 1) * 32 + d.getDayOfMonth();
                    flags[k] |= 4;
                } catch (DateTimeParseException e) { /* similarity 0, Date.java:62-65 */ }
            }
        }
        final int[] s = Arrays.stream(source).map(item::get).toArray(), t = Arrays.stream(target).map(item::get).toArray();
        final float[][] sim = new float[1][];
        final int[][] vo = GraphEmb.similarityPairs(ctx, geMethod, group.upperTriangle, group.threshold, cfg.getSmooth(),
                cfg.getDistance(), (cfg.getTimeEnum().ordinal() + 1) % 3 /* BACKWARDS, FORWARDS, BIDIRECTIONAL -> 1, 2, 0 */,
                vert, stringId, flags, profOff, profKey, profCnt, nCut, cutSlot, numVal, numStatus, epochDay, packed, s, t, charOff, chars, sim);
        final Result r = new Result(); r.vert = vo[0]; r.other = vo[1]; r.similarity = sim[0];
        return r;
    }
}
