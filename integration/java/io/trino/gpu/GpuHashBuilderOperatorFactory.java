/*
 * OperatorFactory for GpuHashBuilderOperator with HashBuilderOperatorFactory's arguments (OP/HashBuilderOperator.java:62-170) minus
 * the spill and filter-function parts.  One build driver per lookup source: partitioned builds (PartitionedLookupSourceFactory,
 * OP/PartitionedLookupSourceFactory.java:110-296) are replaced by one device table per GPU, the rows routed by the exchange.
 * Not compiled in this repository (no JDK).
 */
package io.trino.gpu;

import com.google.common.util.concurrent.SettableFuture;
import io.trino.operator.DriverContext;
import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.operator.OperatorFactory;
import io.trino.spi.type.Type;
import io.trino.sql.planner.plan.PlanNodeId;

import java.util.List;
import java.util.OptionalInt;

public class GpuHashBuilderOperatorFactory
        implements OperatorFactory
{
    private final int operatorId;
    private final PlanNodeId planNodeId;
    private final long ctx;
    private final List<Type> buildTypes;
    private final List<Integer> hashChannels;
    private final OptionalInt preComputedHashChannel;
    private final int expectedPositions;
    private final SettableFuture<GpuLookupSource> lookupSourceFuture = SettableFuture.create();
    private boolean created;

    public GpuHashBuilderOperatorFactory(int operatorId, PlanNodeId planNodeId, long ctx, List<Type> buildTypes, List<Integer> hashChannels,
            OptionalInt preComputedHashChannel, int expectedPositions)
    {
        this.operatorId = operatorId;
        this.planNodeId = planNodeId;
        this.ctx = ctx;
        this.buildTypes = List.copyOf(buildTypes);
        this.hashChannels = List.copyOf(hashChannels);
        this.preComputedHashChannel = preComputedHashChannel;
        this.expectedPositions = expectedPositions;
    }

    /** What the probe side waits on (LookupSourceFactory.createLookupSourceProvider's role, JoinBridgeManager). */
    public SettableFuture<GpuLookupSource> getLookupSourceFuture()
    {
        return lookupSourceFuture;
    }

    public List<Type> getBuildTypes()
    {
        return buildTypes;
    }

    @Override
    public Operator createOperator(DriverContext driverContext)
    {
        if (created) {
            throw new IllegalStateException("the GPU build side runs one driver per lookup source");
        }
        created = true;
        OperatorContext operatorContext = driverContext.addOperatorContext(operatorId, planNodeId, GpuHashBuilderOperator.class.getSimpleName());
        int[] joinChannels = hashChannels.stream().mapToInt(Integer::intValue).toArray();
        return new GpuHashBuilderOperator(operatorContext, ctx, buildTypes, joinChannels, preComputedHashChannel.orElse(-1), expectedPositions, lookupSourceFuture);
    }

    @Override
    public void noMoreOperators() {}

    @Override
    public OperatorFactory duplicate()
    {
        throw new UnsupportedOperationException("the GPU build side cannot be duplicated");
    }
}
