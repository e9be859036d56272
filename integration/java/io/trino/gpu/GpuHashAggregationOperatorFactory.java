/*
 * OperatorFactory for GpuHashAggregationOperator, with the constructor arguments of HashAggregationOperatorFactory
 * (OP/HashAggregationOperator.java:54-262).  LocalExecutionPlanner.createHashAggregationOperatorFactory
 * (MAIN/sql/planner/LocalExecutionPlanner.java:3035-3054) returns this one instead when GpuSupport.supports(...) says every key type
 * and aggregation function maps onto the library (count / sum / avg / min / max over the types of PageArrays.typeTag).
 * Not compiled in this repository (no JDK).
 */
package io.trino.gpu;

import io.airlift.units.DataSize;
import io.trino.operator.DriverContext;
import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.operator.OperatorFactory;
import io.trino.spi.type.Type;
import io.trino.sql.planner.plan.AggregationNode.Step;
import io.trino.sql.planner.plan.PlanNodeId;

import java.util.List;
import java.util.Optional;

public class GpuHashAggregationOperatorFactory
        implements OperatorFactory
{
    /** One aggregate of the plan, already resolved to the library's vocabulary. */
    public static final class GpuAggregate
    {
        final int kind;             // PrestoGpu.AGG_*
        final int inputType;        // type tag of the argument, 0 for count(*)
        final int inputChannel;     // argument channel (raw steps) or first intermediate channel (FINAL / INTERMEDIATE); -1 for count(*)
        final int maskChannel;      // -1: none

        public GpuAggregate(int kind, int inputType, int inputChannel, int maskChannel)
        {
            this.kind = kind;
            this.inputType = inputType;
            this.inputChannel = inputChannel;
            this.maskChannel = maskChannel;
        }
    }

    private final int operatorId;
    private final PlanNodeId planNodeId;
    private final long ctx;
    private final List<Type> inputTypes;
    private final List<Type> groupByTypes;
    private final List<Integer> groupByChannels;
    private final Step step;
    private final List<GpuAggregate> aggregates;
    private final Optional<Integer> hashChannel;
    private final int expectedGroups;
    private final Optional<DataSize> maxPartialMemory;
    private final List<Type> outputTypes;
    private boolean closed;

    public GpuHashAggregationOperatorFactory(int operatorId, PlanNodeId planNodeId, long ctx, List<Type> inputTypes, List<? extends Type> groupByTypes,
            List<Integer> groupByChannels, Step step, List<GpuAggregate> aggregates, Optional<Integer> hashChannel, int expectedGroups,
            Optional<DataSize> maxPartialMemory, List<Type> outputTypes)
    {
        this.operatorId = operatorId;
        this.planNodeId = planNodeId;
        this.ctx = ctx;
        this.inputTypes = List.copyOf(inputTypes);
        this.groupByTypes = List.copyOf(groupByTypes);
        this.groupByChannels = List.copyOf(groupByChannels);
        this.step = step;
        this.aggregates = List.copyOf(aggregates);
        this.hashChannel = hashChannel;
        this.expectedGroups = expectedGroups;
        this.maxPartialMemory = maxPartialMemory;
        this.outputTypes = List.copyOf(outputTypes);
    }

    static int stepTag(Step step)
    {
        switch (step) {
            case PARTIAL:
                return PrestoGpu.STEP_PARTIAL;
            case FINAL:
                return PrestoGpu.STEP_FINAL;
            case INTERMEDIATE:
                return PrestoGpu.STEP_INTERMEDIATE;
            default:
                return PrestoGpu.STEP_SINGLE;
        }
    }

    @Override
    public Operator createOperator(DriverContext driverContext)
    {
        if (closed) {
            throw new IllegalStateException("Factory is already closed");
        }
        OperatorContext operatorContext = driverContext.addOperatorContext(operatorId, planNodeId, GpuHashAggregationOperator.class.getSimpleName());
        int[] keyChannels = groupByChannels.stream().mapToInt(Integer::intValue).toArray();
        int[] specs = new int[4 * aggregates.size()];
        for (int i = 0; i < aggregates.size(); i++) {
            GpuAggregate a = aggregates.get(i);
            specs[4 * i] = a.kind;
            specs[4 * i + 1] = a.inputType;
            specs[4 * i + 2] = a.inputChannel;
            specs[4 * i + 3] = a.maskChannel;
        }
        long partialBytes = maxPartialMemory.map(DataSize::toBytes).orElse(Long.MAX_VALUE);
        return new GpuHashAggregationOperator(operatorContext, ctx, inputTypes, keyChannels, PageArrays.typeTags(groupByTypes), hashChannel.orElse(-1), specs,
                stepTag(step), expectedGroups, partialBytes, PageArrays.typeTags(outputTypes));
    }

    @Override
    public void noMoreOperators()
    {
        closed = true;
    }

    @Override
    public OperatorFactory duplicate()
    {
        return new GpuHashAggregationOperatorFactory(operatorId, planNodeId, ctx, inputTypes, groupByTypes, groupByChannels, step, aggregates, hashChannel,
                expectedGroups, maxPartialMemory, outputTypes);
    }
}
