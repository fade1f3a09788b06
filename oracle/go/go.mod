module llmconsensus/oracle

go 1.21
