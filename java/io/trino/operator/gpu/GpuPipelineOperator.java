/*
 * The drop-in: an io.trino.operator.Operator (core/trino-main/src/main/java/io/trino/operator/Operator.java:21-103) that
 * runs filter -> project -> (hash-)aggregate on a B200 through libprestogpu. It replaces, in one operator,
 *   FilterAndProjectOperator (FilterAndProjectOperator.java:42-88) + AggregationOperator (AggregationOperator.java:93-186)
 *   | HashAggregationOperator (HashAggregationOperator.java:367-518).
 * NOT COMPILED IN THIS REPOSITORY'S IMAGE (no JDK, no Maven repository). The protocol below is the one the Python mirror
 * (presto_b200/operators.py) and the C ABI tests exercise.
 */
package io.trino.operator.gpu;

import com.google.common.util.concurrent.ListenableFuture;
import com.google.common.util.concurrent.SettableFuture;
import io.trino.operator.DriverContext;
import io.trino.operator.Operator;
import io.trino.operator.OperatorContext;
import io.trino.operator.OperatorFactory;
import io.trino.spi.Page;
import io.trino.spi.block.Block;
import io.trino.spi.type.Type;
import io.trino.sql.planner.plan.PlanNodeId;

import java.util.List;

import static com.google.common.base.Preconditions.checkState;
import static java.util.Objects.requireNonNull;

public class GpuPipelineOperator
        implements Operator
{
    public static class GpuPipelineOperatorFactory
            implements OperatorFactory
    {
        private final int operatorId;
        private final PlanNodeId planNodeId;
        private final byte[] plan;          // plan text produced by GpuPlanTranslator from RowExpressions + aggregation spec
        private final List<Type> inputTypes;
        private final List<Type> outputTypes;
        private final int device;
        private boolean closed;

        public GpuPipelineOperatorFactory(int operatorId, PlanNodeId planNodeId, byte[] plan, List<Type> inputTypes, List<Type> outputTypes, int device)
        {
            this.inputTypes = requireNonNull(inputTypes, "inputTypes is null");
            this.operatorId = operatorId;
            this.planNodeId = requireNonNull(planNodeId, "planNodeId is null");
            this.plan = requireNonNull(plan, "plan is null");
            this.outputTypes = requireNonNull(outputTypes, "outputTypes is null");
            this.device = device;
            // plan-time rejection: compile-only create (device = -1) throws TrinoException(NOT_SUPPORTED) for anything
            // outside the GPU subset, exactly where LocalExecutionPlanner would otherwise build the CPU factory
            PrestoGpuNative.destroy(PrestoGpuNative.create(plan, -1));
        }

        @Override
        public Operator createOperator(DriverContext driverContext)
        {
            checkState(!closed, "Factory is already closed");
            OperatorContext operatorContext = driverContext.addOperatorContext(operatorId, planNodeId, GpuPipelineOperator.class.getSimpleName());
            return new GpuPipelineOperator(operatorContext, PrestoGpuNative.create(plan, device), inputTypes, outputTypes);
        }

        @Override
        public void noMoreOperators()
        {
            closed = true;
        }

        @Override
        public OperatorFactory duplicate()
        {
            return new GpuPipelineOperatorFactory(operatorId, planNodeId, plan, inputTypes, outputTypes, device);
        }
    }

    private final OperatorContext operatorContext;
    private final List<Type> inputTypes;
    private final List<Type> outputTypes;
    private long handle;

    private GpuPipelineOperator(OperatorContext operatorContext, long handle, List<Type> inputTypes, List<Type> outputTypes)
    {
        this.operatorContext = requireNonNull(operatorContext, "operatorContext is null");
        this.handle = handle;
        this.inputTypes = inputTypes;
        this.outputTypes = outputTypes;
    }

    /**
     * Operator.isBlocked() (Operator.java:24-35): while device work queued by this operator is still running, getOutput() would
     * have to wait for it; instead of blocking the Driver thread the operator hands out a future that a CUDA host callback completes.
     */
    @Override
    public ListenableFuture<Void> isBlocked()
    {
        if (handle == 0 || !PrestoGpuNative.isBlocked(handle)) {
            return NOT_BLOCKED;
        }
        SettableFuture<Void> done = SettableFuture.create();
        PrestoGpuNative.onUnblocked(handle, () -> done.set(null));
        return done;
    }

    @Override
    public OperatorContext getOperatorContext()
    {
        return operatorContext;
    }

    @Override
    public boolean needsInput()
    {
        return PrestoGpuNative.needsInput(handle);
    }

    @Override
    public void addInput(Page page)
    {
        checkState(needsInput(), "Operator does not need input");
        GpuPageMarshaller.addPage(handle, page, inputTypes);    // getLoadedBlock() per channel, raw arrays -> PrestoGpuNative.addPage
    }

    @Override
    public Page getOutput()
    {
        Object[] raw = PrestoGpuNative.getOutput(handle);
        if (raw == null) {
            return null;
        }
        Block[] blocks = GpuPageMarshaller.toBlocks(raw, outputTypes);   // fresh LongArrayBlock / VariableWidthBlock ...
        return new Page((Integer) raw[0], blocks);
    }

    @Override
    public void finish()
    {
        PrestoGpuNative.finish(handle);
    }

    @Override
    public boolean isFinished()
    {
        return PrestoGpuNative.isFinished(handle);
    }

    @Override
    public void close()
    {
        if (handle != 0) {
            PrestoGpuNative.destroy(handle);
            handle = 0;
        }
    }
}
