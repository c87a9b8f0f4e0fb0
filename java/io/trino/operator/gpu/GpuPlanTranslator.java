/*
 * RowExpressions + an aggregation spec -> the plan text libprestogpu compiles (INTEGRATION.md §3, presto_b200/csrc/plan.cpp).
 * The Java-side twin of presto_b200/relational.py + operators.plan_text, which the test-suite exercises; what it emits for the
 * benchmark queries is byte-for-byte what tests/ feed to pgpu_pipeline_create.
 *
 * Everything outside the supported subset raises TrinoException(NOT_SUPPORTED) HERE, at plan time, so that
 * LocalExecutionPlanner can keep the CPU operators for that plan node (there is no fallback inside the GPU operator).
 * Reference types used: io.trino.sql.relational.{RowExpression, RowExpressionVisitor, CallExpression, SpecialForm,
 * ConstantExpression, InputReferenceExpression} (core/trino-main/src/main/java/io/trino/sql/relational/),
 * io.trino.metadata.Signature.{isOperatorName, unmangleOperator} (Signature.java:77-92), io.trino.spi.type.*.
 * NOT COMPILED IN THIS REPOSITORY'S IMAGE (no JDK, no Maven repository).
 */
package io.trino.operator.gpu;

import io.airlift.slice.Slice;
import io.trino.spi.TrinoException;
import io.trino.spi.function.OperatorType;
import io.trino.spi.type.DecimalType;
import io.trino.spi.type.Decimals;
import io.trino.spi.type.Type;
import io.trino.spi.type.VarcharType;
import io.trino.sql.relational.CallExpression;
import io.trino.sql.relational.ConstantExpression;
import io.trino.sql.relational.InputReferenceExpression;
import io.trino.sql.relational.LambdaDefinitionExpression;
import io.trino.sql.relational.RowExpression;
import io.trino.sql.relational.RowExpressionVisitor;
import io.trino.sql.relational.SpecialForm;
import io.trino.sql.relational.VariableReferenceExpression;

import java.nio.charset.StandardCharsets;
import java.util.List;
import java.util.Optional;
import java.util.StringJoiner;

import static io.trino.metadata.Signature.isOperatorName;
import static io.trino.metadata.Signature.unmangleOperator;
import static io.trino.spi.StandardErrorCode.NOT_SUPPORTED;
import static io.trino.spi.type.BigintType.BIGINT;
import static io.trino.spi.type.BooleanType.BOOLEAN;
import static io.trino.spi.type.DateType.DATE;
import static io.trino.spi.type.DoubleType.DOUBLE;
import static io.trino.spi.type.IntegerType.INTEGER;
import static io.trino.spi.type.VarbinaryType.VARBINARY;

public final class GpuPlanTranslator
{
    private GpuPlanTranslator() {}

    /** One accumulator binding: what AccumulatorFactory carries after InternalAggregationFunction.bind(channels, mask). */
    public static final class Aggregate
    {
        final String function;            // sum | count | avg | min | max
        final Type outputType;
        final Optional<Integer> argument; // channel of the projected page (first intermediate channel for FINAL / INTERMEDIATE)
        final Type argumentType;          // raw argument type (also for FINAL: it selects the state layout)
        final Optional<Integer> mask;

        public Aggregate(String function, Type outputType, Optional<Integer> argument, Type argumentType, Optional<Integer> mask)
        {
            this.function = function;
            this.outputType = outputType;
            this.argument = argument;
            this.argumentType = argumentType;
            this.mask = mask;
        }
    }

    /** step: single | partial | intermediate | final (AggregationNode.Step, lower case). */
    public static byte[] translate(
            List<Type> inputTypes,
            Optional<RowExpression> filter,
            List<RowExpression> projections,
            Optional<String> step,
            List<Integer> groupByChannels,
            List<Aggregate> aggregates,
            int expectedGroups)
    {
        StringBuilder plan = new StringBuilder("(plan\n (input");
        for (Type type : inputTypes) {
            plan.append(' ').append(type(type));
        }
        plan.append(")\n");
        Emitter emitter = new Emitter();
        filter.ifPresent(expression -> plan.append(" (filter ").append(expression.accept(emitter, null)).append(")\n"));
        if (!projections.isEmpty()) {
            StringJoiner joiner = new StringJoiner(" ", " (project ", ")\n");
            projections.forEach(projection -> joiner.add(projection.accept(emitter, null)));
            plan.append(joiner);
        }
        if (step.isPresent()) {
            plan.append(" (aggregate (step ").append(step.get()).append(")\n");
            if (!groupByChannels.isEmpty()) {
                StringJoiner group = new StringJoiner(" ", " (group ", ")\n");
                groupByChannels.forEach(channel -> group.add(Integer.toString(channel)));
                plan.append(group);
            }
            plan.append(" (expected-groups ").append(expectedGroups).append(")\n");
            for (Aggregate aggregate : aggregates) {
                plan.append(" (agg ").append(aggregate.function).append(' ').append(type(aggregate.outputType));
                aggregate.argument.ifPresent(channel -> plan.append(" (arg ").append(channel).append(' ').append(type(aggregate.argumentType)).append(')'));
                aggregate.mask.ifPresent(channel -> plan.append(" (mask ").append(channel).append(')'));
                plan.append(")\n");
            }
            plan.append(")\n");
        }
        return plan.append(")").toString().getBytes(StandardCharsets.US_ASCII);
    }

    /**
     * Filter + project only, with the thresholds FilterAndProjectOperator (main/operator/FilterAndProjectOperator.java:75-83) hands to
     * MergePages: minOutputPageSize (session property filter_and_project_min_output_page_size) and minOutputPageRowCount.
     */
    public static byte[] translateFilterAndProject(
            List<Type> inputTypes,
            Optional<RowExpression> filter,
            List<RowExpression> projections,
            long minOutputPageSizeInBytes,
            int minOutputPageRowCount)
    {
        String plan = new String(translate(inputTypes, filter, projections, Optional.empty(), List.of(), List.of(), 0), StandardCharsets.US_ASCII);
        return (plan.substring(0, plan.length() - 1)
                + " (options (merge-min-rows " + minOutputPageRowCount + ") (merge-min-bytes " + minOutputPageSizeInBytes + ")))")
                .getBytes(StandardCharsets.US_ASCII);
    }

    /** Type names of the plan language (csrc/plan.cpp parseType). */
    static String type(Type type)
    {
        if (type.equals(BOOLEAN)) {
            return "boolean";
        }
        if (type.equals(BIGINT)) {
            return "bigint";
        }
        if (type.equals(INTEGER)) {
            return "integer";
        }
        if (type.equals(DATE)) {
            return "date";
        }
        if (type.equals(DOUBLE)) {
            return "double";
        }
        if (type.equals(VARBINARY)) {
            return "varbinary";
        }
        if (type instanceof DecimalType) {
            DecimalType decimal = (DecimalType) type;
            return "(decimal " + decimal.getPrecision() + " " + decimal.getScale() + ")";
        }
        if (type instanceof VarcharType) {
            VarcharType varchar = (VarcharType) type;
            return varchar.isUnbounded() ? "varchar" : "(varchar " + varchar.getBoundedLength() + ")";
        }
        throw unsupported("type " + type.getDisplayName());
    }

    static TrinoException unsupported(String what)
    {
        return new TrinoException(NOT_SUPPORTED, "not supported on the GPU path: " + what);
    }

    /** RowExpression -> S-expression. Mirrors RowExpressionCompiler's dispatch (main/sql/gen/RowExpressionCompiler.java:95-190). */
    private static final class Emitter
            implements RowExpressionVisitor<String, Void>
    {
        @Override
        public String visitInputReference(InputReferenceExpression reference, Void context)
        {
            return "(in " + reference.getField() + " " + type(reference.getType()) + ")";
        }

        @Override
        public String visitConstant(ConstantExpression literal, Void context)
        {
            Type type = literal.getType();
            Object value = literal.getValue();
            String text;
            if (value == null) {
                text = "null";
            }
            else if (type.equals(BOOLEAN)) {
                text = ((Boolean) value) ? "true" : "false";
            }
            else if (type.equals(DOUBLE)) {
                text = String.format("0x%016x", Double.doubleToRawLongBits((Double) value));   // raw IEEE bits: no decimal round trip
            }
            else if (type.equals(BIGINT) || type.equals(INTEGER) || type.equals(DATE)) {
                text = Long.toString((Long) value);
            }
            else if (type instanceof DecimalType) {
                // unscaled value: a long for short decimals, the sign-magnitude Slice for long ones (Decimals.java:184-190)
                text = Decimals.isShortDecimal(type) ? Long.toString((Long) value) : Decimals.decodeUnscaledValue((Slice) value).toString();
            }
            else {
                throw unsupported("constant of type " + type.getDisplayName());
            }
            return "(const " + type(type) + " " + text + ")";
        }

        @Override
        public String visitCall(CallExpression call, Void context)
        {
            String name = call.getResolvedFunction().getSignature().getName();
            String operator;
            if (isOperatorName(name)) {
                operator = operatorName(unmangleOperator(name));
            }
            else if (name.equalsIgnoreCase("not")) {   // SqlToRowExpressionTranslator.java:311 resolves NOT as the function "not"
                operator = "not";
            }
            else {
                throw unsupported("scalar function '" + name + "' (only operators, NOT and CAST compile to the GPU)");
            }
            return "(call " + operator + " " + type(call.getType()) + arguments(call.getArguments()) + ")";
        }

        // The analyzer only emits EQUAL, LESS_THAN and LESS_THAN_OR_EQUAL (a > b arrives as b < a, a <> b as not(a = b)):
        // io.trino.spi.function.OperatorType has no GREATER_THAN / NOT_EQUAL. The plan language also accepts gt / ge / ne
        // for hand-written plans.
        private static String operatorName(OperatorType operator)
        {
            switch (operator) {
                case ADD:
                    return "add";
                case SUBTRACT:
                    return "subtract";
                case MULTIPLY:
                    return "multiply";
                case DIVIDE:
                    return "divide";
                case MODULUS:
                    return "modulus";
                case NEGATION:
                    return "negate";
                case EQUAL:
                    return "eq";
                case LESS_THAN:
                    return "lt";
                case LESS_THAN_OR_EQUAL:
                    return "le";
                case CAST:
                    return "cast";
                default:
                    throw unsupported("operator " + operator);
            }
        }

        @Override
        public String visitSpecialForm(SpecialForm specialForm, Void context)
        {
            String arguments = arguments(specialForm.getArguments());
            switch (specialForm.getForm()) {
                case AND:
                    return "(and" + arguments + ")";
                case OR:
                    return "(or" + arguments + ")";
                case BETWEEN:
                    return "(between" + arguments + ")";
                case IN:
                    return "(in-list" + arguments + ")";
                case IS_NULL:
                    return "(is-null" + arguments + ")";
                case IF:
                    return "(if " + type(specialForm.getType()) + arguments + ")";
                case COALESCE:
                    return "(coalesce " + type(specialForm.getType()) + arguments + ")";
                case NULL_IF:
                    // the casts to the comparison's common type travel as function dependencies (NullIfCodeGenerator.java:57-60):
                    // only the same-type form is expressible
                    if (!specialForm.getArguments().get(0).getType().equals(specialForm.getArguments().get(1).getType())) {
                        throw unsupported("NULLIF over operands of different types");
                    }
                    return "(null-if" + arguments + ")";
                case SWITCH:
                    return "(switch " + type(specialForm.getType()) + arguments + ")";
                case WHEN:
                    return "(when" + arguments + ")";
                default:
                    throw unsupported("special form " + specialForm.getForm());
            }
        }

        @Override
        public String visitLambda(LambdaDefinitionExpression lambda, Void context)
        {
            throw unsupported("lambda expressions");
        }

        @Override
        public String visitVariableReference(VariableReferenceExpression reference, Void context)
        {
            throw unsupported("variable references");
        }

        private String arguments(List<RowExpression> arguments)
        {
            StringBuilder text = new StringBuilder();
            for (RowExpression argument : arguments) {
                text.append(' ').append(argument.accept(this, null));
            }
            return text.toString();
        }
    }
}
