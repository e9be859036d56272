/* Caller-allocated output arrays for pg_hashagg_get_output / pg_lookup_gather (one set per output channel). */
package io.trino.gpu;

final class OutputArrays
{
    Object[] values;
    Object[] nulls;
    Object[] offsets;
    long[] valueBytes;

    static OutputArrays allocate(int[] types, int rows)
    {
        return allocate(types, rows, new long[types.length]);
    }

    /** varWidthBytes[c] comes from pg_hashagg_varwidth_bytes / pg_lookup_gather_var_bytes for VARCHAR/VARBINARY channels. */
    static OutputArrays allocate(int[] types, int rows, long[] varWidthBytes)
    {
        OutputArrays out = new OutputArrays();
        int n = types.length;
        out.values = new Object[n];
        out.nulls = new Object[n];
        out.offsets = new Object[n];
        out.valueBytes = varWidthBytes;
        for (int c = 0; c < n; c++) {
            out.nulls[c] = new boolean[rows];
            switch (types[c]) {
                case PrestoGpu.INTEGER:
                case PrestoGpu.DATE:
                    out.values[c] = new int[rows];
                    break;
                case PrestoGpu.LONG_DECIMAL:
                    out.values[c] = new long[2 * rows];
                    break;
                case PrestoGpu.BOOLEAN:
                    out.values[c] = new byte[rows];
                    break;
                case PrestoGpu.VARCHAR:
                case PrestoGpu.VARBINARY:
                    out.values[c] = new byte[(int) Math.max(1, varWidthBytes[c])];
                    out.offsets[c] = new int[rows + 1];
                    break;
                default:
                    out.values[c] = new long[rows];
            }
        }
        return out;
    }
}
