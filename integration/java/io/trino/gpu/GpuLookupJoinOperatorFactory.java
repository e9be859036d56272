/*
 * OperatorFactory for GpuLookupJoinOperator (LookupJoinOperators.innerJoin / probeOuterJoin -> LookupJoinOperatorFactory,
 * OP/LookupJoinOperators.java:36-82, OP/LookupJoinOperatorFactory.java).  Probe operators share the build side's GpuLookupSource;
 * the last one to close releases the device table (the job of JoinBridgeManager / ReferenceCount in the reference).
 * Not compiled in this repository (no JDK).
 */
package io.trino.gpu;

import io.trino.operator.DriverContext;
import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.operator.OperatorFactory;
import io.trino.spi.type.Type;
import io.trino.sql.planner.plan.PlanNodeId;

import java.util.List;
import java.util.Optional;
import java.util.concurrent.atomic.AtomicInteger;

public class GpuLookupJoinOperatorFactory
        implements OperatorFactory
{
    private final int operatorId;
    private final PlanNodeId planNodeId;
    private final GpuHashBuilderOperatorFactory buildSide;
    private final List<Type> probeTypes;
    private final List<Integer> probeJoinChannels;
    private final Optional<Integer> probeHashChannel;
    private final List<Integer> probeOutputChannels;
    private final List<Integer> buildOutputChannels;
    private final int joinType;
    private final AtomicInteger users = new AtomicInteger(1);      // the factory itself, until noMoreOperators()

    public GpuLookupJoinOperatorFactory(int operatorId, PlanNodeId planNodeId, GpuHashBuilderOperatorFactory buildSide, List<Type> probeTypes,
            List<Integer> probeJoinChannels, Optional<Integer> probeHashChannel, List<Integer> probeOutputChannels, List<Integer> buildOutputChannels, int joinType)
    {
        this.operatorId = operatorId;
        this.planNodeId = planNodeId;
        this.buildSide = buildSide;
        this.probeTypes = List.copyOf(probeTypes);
        this.probeJoinChannels = List.copyOf(probeJoinChannels);
        this.probeHashChannel = probeHashChannel;
        this.probeOutputChannels = List.copyOf(probeOutputChannels);
        this.buildOutputChannels = List.copyOf(buildOutputChannels);
        this.joinType = joinType;
    }

    private void release()
    {
        if (users.decrementAndGet() == 0 && buildSide.getLookupSourceFuture().isDone()) {
            try {
                buildSide.getLookupSourceFuture().get().close();
            }
            catch (Exception ignored) {
                // the build side failed: nothing to free
            }
        }
    }

    private static int[] toArray(List<Integer> list)
    {
        return list.stream().mapToInt(Integer::intValue).toArray();
    }

    @Override
    public Operator createOperator(DriverContext driverContext)
    {
        OperatorContext operatorContext = driverContext.addOperatorContext(operatorId, planNodeId, GpuLookupJoinOperator.class.getSimpleName());
        users.incrementAndGet();
        List<Type> joinTypes = probeJoinChannels.stream().map(probeTypes::get).collect(java.util.stream.Collectors.toList());
        return new GpuLookupJoinOperator(operatorContext, buildSide.getLookupSourceFuture(), this::release, toArray(probeJoinChannels), joinTypes, probeHashChannel,
                toArray(probeOutputChannels), toArray(buildOutputChannels), joinType);
    }

    @Override
    public void noMoreOperators()
    {
        release();
    }

    @Override
    public OperatorFactory duplicate()
    {
        return new GpuLookupJoinOperatorFactory(operatorId, planNodeId, buildSide, probeTypes, probeJoinChannels, probeHashChannel, probeOutputChannels,
                buildOutputChannels, joinType);
    }
}
